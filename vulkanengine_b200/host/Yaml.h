// Reader for the YAML subset the reference's serializers emit through yaml-cpp's YAML::Emitter
// (VulkanCore/vendor/yaml-cpp; SceneSerializer.cpp:163-320, AssetManager.cpp:67-93, MaterialAssetImporter.cpp):
// block maps, block sequences of maps ("- Key: value"), flow sequences of scalars ("[a, b, c]"), plain and quoted
// scalars.  Anchors, tags, multi-line scalars and multi-document streams do not occur in those files and are not
// handled.  Header-only; used by SceneSerializer.cpp / AssetManager (host mirror).
#pragma once
#include <cstdint>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace VulkanCore::Yaml {

class Node {
public:
    enum class Kind { Null, Scalar, Map, Seq };
    Kind kind = Kind::Null;
    std::string scalar;
    std::vector<std::pair<std::string, Node>> map;  // insertion order kept, like YAML::Node iteration
    std::vector<Node> seq;

    explicit operator bool() const { return kind != Kind::Null; }
    const Node& operator[](const std::string& key) const
    {
        static const Node null;
        if (kind == Kind::Map)
            for (const auto& kv : map)
                if (kv.first == key) return kv.second;
        return null;
    }
    const Node& operator[](size_t i) const { return seq.at(i); }
    size_t size() const { return kind == Kind::Seq ? seq.size() : kind == Kind::Map ? map.size() : 0; }

    const std::string& str() const
    {
        if (kind != Kind::Scalar) throw std::runtime_error("yaml: scalar expected");
        return scalar;
    }
    uint64_t u64() const { return strtoull(str().c_str(), nullptr, 10); }
    int64_t i64() const { return strtoll(str().c_str(), nullptr, 10); }
    float f32() const { return strtof(str().c_str(), nullptr); }
    double f64() const { return strtod(str().c_str(), nullptr); }
    bool boolean() const { return scalar == "true" || scalar == "True" || scalar == "1"; }
};

namespace detail {

struct Line {
    int indent;
    std::string text;  // without indentation and without a trailing comment
};

inline std::string trim(const std::string& s)
{
    size_t a = 0, b = s.size();
    while (a < b && (s[a] == ' ' || s[a] == '\t' || s[a] == '\r')) ++a;
    while (b > a && (s[b - 1] == ' ' || s[b - 1] == '\t' || s[b - 1] == '\r')) --b;
    return s.substr(a, b - a);
}

inline Node scalar_node(std::string v)
{
    Node n;
    v = trim(v);
    if (v == "~" || v == "null") return n;
    if (v.size() >= 2 && ((v.front() == '"' && v.back() == '"') || (v.front() == '\'' && v.back() == '\''))) v = v.substr(1, v.size() - 2);
    n.kind = Node::Kind::Scalar;
    n.scalar = v;
    return n;
}

inline Node value_node(const std::string& raw)
{
    const std::string v = trim(raw);
    if (v == "{}") {
        Node n;
        n.kind = Node::Kind::Map;
        return n;
    }
    if (!v.empty() && v.front() == '[') {
        if (v.back() != ']') throw std::runtime_error("yaml: unterminated flow sequence");
        Node n;
        n.kind = Node::Kind::Seq;
        const std::string body = v.substr(1, v.size() - 2);
        size_t start = 0;
        while (start <= body.size()) {
            size_t comma = body.find(',', start);
            if (comma == std::string::npos) comma = body.size();
            const std::string item = trim(body.substr(start, comma - start));
            if (!item.empty()) n.seq.push_back(scalar_node(item));
            start = comma + 1;
        }
        return n;
    }
    return scalar_node(v);
}

// splits "key: value" / "key:" at the first ": " (or trailing ':') outside quotes; false = not a map entry
inline bool split_key(const std::string& t, std::string& key, std::string& rest)
{
    if (!t.empty() && (t.front() == '[' || t.front() == '{')) return false;
    bool quoted = false;
    for (size_t i = 0; i < t.size(); ++i) {
        if (t[i] == '"') quoted = !quoted;
        if (!quoted && t[i] == ':' && (i + 1 == t.size() || t[i + 1] == ' ')) {
            key = trim(t.substr(0, i));
            if (key.size() >= 2 && key.front() == '"' && key.back() == '"') key = key.substr(1, key.size() - 2);
            rest = i + 1 < t.size() ? trim(t.substr(i + 1)) : std::string();
            return true;
        }
    }
    return false;
}

struct Parser {
    std::vector<Line> lines;
    size_t pos = 0;

    Node block(int indent)
    {
        if (pos >= lines.size() || lines[pos].indent < indent) return Node();
        const int ind = lines[pos].indent;
        const std::string& first = lines[pos].text;
        if (first == "-" || first.rfind("- ", 0) == 0) return sequence(ind);
        std::string k, r;
        if (split_key(first, k, r)) return mapping(ind);
        Node n = value_node(first);
        ++pos;
        return n;
    }

    Node sequence(int ind)
    {
        Node n;
        n.kind = Node::Kind::Seq;
        while (pos < lines.size() && lines[pos].indent == ind && (lines[pos].text == "-" || lines[pos].text.rfind("- ", 0) == 0)) {
            // the item's content starts two columns further in; re-read the same line as the item's first line
            std::string rest = lines[pos].text.size() > 2 ? trim(lines[pos].text.substr(2)) : std::string();
            if (rest.empty()) {
                ++pos;
                n.seq.push_back(block(ind + 1));
            } else {
                lines[pos].indent = ind + 2;
                lines[pos].text = rest;
                n.seq.push_back(block(ind + 2));
            }
        }
        return n;
    }

    Node mapping(int ind)
    {
        Node n;
        n.kind = Node::Kind::Map;
        while (pos < lines.size() && lines[pos].indent == ind) {
            std::string key, rest;
            if (!split_key(lines[pos].text, key, rest)) throw std::runtime_error("yaml: map entry expected: " + lines[pos].text);
            ++pos;
            if (!rest.empty()) {
                n.map.emplace_back(key, value_node(rest));
            } else if (pos < lines.size() && lines[pos].indent > ind) {
                n.map.emplace_back(key, block(lines[pos].indent));
            } else if (pos < lines.size() && lines[pos].indent == ind && (lines[pos].text == "-" || lines[pos].text.rfind("- ", 0) == 0)) {
                n.map.emplace_back(key, sequence(ind));  // a sequence may sit at its key's own indentation
            } else {
                n.map.emplace_back(key, Node());
            }
        }
        return n;
    }
};

}  // namespace detail

inline Node Load(const std::string& text)
{
    detail::Parser p;
    size_t start = 0;
    while (start <= text.size()) {
        size_t nl = text.find('\n', start);
        if (nl == std::string::npos) nl = text.size();
        std::string raw = text.substr(start, nl - start);
        start = nl + 1;
        int indent = 0;
        while ((size_t)indent < raw.size() && raw[indent] == ' ') ++indent;
        std::string t = detail::trim(raw);
        if (t.empty() || t[0] == '#' || t == "---" || t == "...") continue;
        p.lines.push_back({indent, t});
    }
    return p.block(0);
}

}  // namespace VulkanCore::Yaml
