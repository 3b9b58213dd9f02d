// Headless C++ host mirror of the reference's scene module — math types and the ECS registry.
//
// The reference keeps its per-frame submission behind C++ member functions; this
// directory re-creates just enough of that host side — same class and method names, same argument
// meaning, same early-outs — to call libscenesubmit through its C ABI without a Vulkan device.
// Nothing here is copied from the reference or its vendored libraries: the registry below is a
// small sparse-set ECS written for this purpose that reproduces the one entt behaviour the path
// depends on, the iteration order of a multi-component view.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <typeindex>
#include <unordered_map>
#include <vector>

namespace VulkanCore {

struct vec3 {
    float x = 0.f, y = 0.f, z = 0.f;
};

// column-major 4x4, same memory as glm::mat4: c[col][row]
struct mat4 {
    float c[4][4];
    static mat4 identity()
    {
        mat4 m;
        std::memset(&m, 0, sizeof m);
        m.c[0][0] = m.c[1][1] = m.c[2][2] = m.c[3][3] = 1.f;
        return m;
    }
    const float* data() const { return &c[0][0]; }
};

using UUID = uint64_t;        // VulkanCore/src/VulkanCore/Core/UUID.h:5-16
using AssetHandle = uint64_t;

// ---- sparse-set registry ---------------------------------------------------------------------------
using entity = uint32_t;
constexpr entity null_entity = 0xFFFFFFFFu;

struct PoolBase {
    std::vector<uint32_t> sparse;  // entity -> position in packed, or ~0u
    std::vector<entity> packed;    // in insertion order (swap-and-pop on erase)
    virtual ~PoolBase() = default;
    bool contains(entity e) const { return e < sparse.size() && sparse[e] != 0xFFFFFFFFu; }
    size_t size() const { return packed.size(); }
};

template <class T>
struct Pool : PoolBase {
    std::vector<T> data;  // parallel to packed: for TransformComponent this IS the 36-byte TRS stream
};

class Registry {
public:
    entity create()
    {
        return m_Next++;
    }

    template <class T, class... Args>
    T& emplace(entity e, Args&&... args)
    {
        Pool<T>& p = pool<T>();
        if (p.sparse.size() <= e) p.sparse.resize(e + 1, 0xFFFFFFFFu);
        p.sparse[e] = (uint32_t)p.packed.size();
        p.packed.push_back(e);
        p.data.emplace_back(std::forward<Args>(args)...);
        return p.data.back();
    }

    template <class T>
    bool all_of(entity e) const
    {
        auto it = m_Pools.find(std::type_index(typeid(T)));
        return it != m_Pools.end() && it->second->contains(e);
    }

    template <class T>
    T& get(entity e)
    {
        Pool<T>& p = pool<T>();
        return p.data[p.sparse[e]];
    }

    template <class T>
    Pool<T>& storage()
    {
        return pool<T>();
    }

    // view<A, B>: like entt, the pool with the fewest elements leads (the first one on a tie) and its packed
    // array is walked from the back to the front; entities missing the other component are skipped.
    // That order is the order of instances inside every draw (SURVEY.md §7 hard part 4).
    template <class A, class B, class F>
    void view(F&& f)
    {
        Pool<A>& pa = pool<A>();
        Pool<B>& pb = pool<B>();
        const bool a_leads = pa.size() <= pb.size();
        const PoolBase& lead = a_leads ? static_cast<PoolBase&>(pa) : static_cast<PoolBase&>(pb);
        for (size_t k = lead.packed.size(); k-- > 0;) {
            const entity e = lead.packed[k];
            if (pa.contains(e) && pb.contains(e)) f(e, pa.data[pa.sparse[e]], pb.data[pb.sparse[e]]);
        }
    }

    size_t alive() const { return m_Next; }

private:
    template <class T>
    Pool<T>& pool()
    {
        auto& slot = m_Pools[std::type_index(typeid(T))];
        if (!slot) slot = std::make_unique<Pool<T>>();
        return static_cast<Pool<T>&>(*slot);
    }
    std::unordered_map<std::type_index, std::unique_ptr<PoolBase>> m_Pools;
    entity m_Next = 0;
};

}  // namespace VulkanCore
