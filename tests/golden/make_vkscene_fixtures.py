"""Extracts the reference's six bundled scenes (VulkanCore/assets/scenes/*.vkscene, AssetRegistry.vkr,
materials/*.vkmat) into tests/golden/vkscenes.json.  Run in the build container, where /root/reference
exists:  python tests/golden/make_vkscene_fixtures.py

What is taken from the reference's files: entity order, UUIDs, TransformComponent values, MeshComponent
handles, material tables (index -> handle), and which materials carry a displacement texture (tessellated
pass).  What is NOT available: submesh tables and vertex bounds, which the reference obtains from the FBX
sources through Assimp (vendored as source with a CMake build: unbuildable here).  Every mesh
therefore gets ONE stand-in submesh with a unit box; the scenes still exercise the real entity lists."""
import glob
import json
import os

import yaml

REF = "/root/reference/VulkanCore/assets"
HERE = os.path.dirname(os.path.abspath(__file__))


def main(REF=REF, out_name="vkscenes.json", source="VulkanCore/assets/scenes/*.vkscene @ reference commit fe134a6c"):
    reg = yaml.safe_load(open(os.path.join(REF, "AssetRegistry.vkr")))["Asset Registry"]
    by_handle = {int(e["Handle"]): e["Metadata"] for e in reg}
    materials = {}
    for h, md in by_handle.items():
        if md["Type"] == "Material":
            path = os.path.join(REF, md["Filepath"])
            tess = False
            if os.path.exists(path):
                m = yaml.safe_load(open(path))["Material"]
                dh = int(m.get("DisplacementHandle", 0) or 0)
                tess = dh != 0 and dh in by_handle  # MaterialAssetImporter.cpp:144-181: the handle must be a valid asset
            materials[str(h)] = {"file": md["Filepath"], "tessellated": tess}
    scenes = {}
    for path in sorted(glob.glob(os.path.join(REF, "scenes", "*.vkscene"))):
        doc = yaml.safe_load(open(path))
        ents = []
        for e in doc["Entities"]:
            # entities are created with a default TransformComponent (Scene::CreateEntityWithUUID); lights are
            # serialized without one
            tc = e.get("TransformComponent", {"Translation": [0, 0, 0], "Rotation": [0, 0, 0], "Scale": [1, 1, 1]})
            ent = {"uuid": str(int(e["Entity"])), "tag": e.get("TagComponent", {}).get("Tag", ""),
                   "has_transform": "TransformComponent" in e,
                   "translation": [float(x) for x in tc["Translation"]], "rotation": [float(x) for x in tc["Rotation"]],
                   "scale": [float(x) for x in tc["Scale"]], "mesh": None, "materials": {}}
            mc = e.get("MeshComponent")
            if mc:
                ent["mesh"] = str(int(mc["MeshHandle"]))
                ent["materials"] = {str(int(k)): str(int(v)) for k, v in (mc.get("MaterialTable") or {}).items()}
            dl = e.get("DirectionalLightComponent")
            if dl and "Direction" in dl:
                ent["light_direction"] = [float(x) for x in dl["Direction"]]
            # the submission walk is view<TransformComponent, MeshComponent>: a mesh without a transform would never draw
            assert ent["has_transform"] or ent["mesh"] is None, (path, ent["uuid"])
            ents.append(ent)
        scenes[os.path.splitext(os.path.basename(path))[0]] = ents
    out = {"source": source, "materials": materials, "scenes": scenes}
    json.dump(out, open(os.path.join(HERE, out_name), "w"), indent=1)
    print({k: (len(v), sum(1 for e in v if e["mesh"])) for k, v in scenes.items()})


if __name__ == "__main__":
    main()
