"""Generates the committed fixtures under tests/golden/ from the REFERENCE'S OWN code (oracle/_ref/libclrt_ref.so: the
reference headers + its vendored tinyobj, compiled in place from /root/reference).  Run in the build container only:

    python tests/golden/make_golden.py

Outputs
  scenes/<name>.npz      packed v0.6 scene buffer of each bundled TestModels mesh + LIGHT 0.0 10.0 -330.0 1000.0
                         (R/Bin/*.txt), packed by the restated Scene::load loop over the reference's tinyobj.
  bvh_<name>.npz         reference outputs for that scene: sorted pairs (canonical stable tie order), nodes (FIXED node
                         count, garbage lanes zeroed) — full arrays for the small meshes, SHA-256 digests for the large —
                         plus primary contacts and host-shaded colours at 96x64 for the demo camera and one rotated pose.
"""
import hashlib
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle_py as O  # noqa: E402
from openclraytracer_b200 import layouts as L  # noqa: E402

MODELS = {
    "cube": "Cube/cube.obj",
    "gardenlight": "Gardenlight/gardenlight01.obj",
    "tenryuu": "Tenryuu/light-cruiser-tenryuu.obj",
    "manfigure": "ManFigure/standard-male-figure.obj",
    "clumsydragon": "ClumsyDragon/clumsy-dragon-vray.obj",
}
LIGHT = [[0.0, 10.0, -330.0, 1000.0]]
W, H = 96, 64
POSE_QUAT = (0.9689124, 0.0, 0.2474040, 0.0)      # ~28.6 deg about Y


def masked_nodes(nodes, n):
    m = nodes.copy()
    m["data"][n:, 3] = 0          # never written by the reference (BVH.h:291-321)
    m["bounds"][n:, 1, 3] = 0     # Q5
    if n == 1:
        m["data"][:, 0] = 0xFFFFFFFF
    return m


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).view(np.uint8).tobytes()).hexdigest()


def main():
    O.build(("ref",))
    for name, rel in MODELS.items():
        scene = O.ref_scene_pack(LIGHT, [os.path.join(O.REFERENCE_ROOT, "TestModels", rel)])
        np.savez_compressed(os.path.join(HERE, "scenes", f"{name}.npz"), scene=scene)
        n = int(np.frombuffer(scene[:96].tobytes(), dtype=L.SCENE_HEADER)[0]["totalNumberOfTriangles"])
        b = O.ref_bvh_build(scene, literal=False)
        nodes = masked_nodes(b["nodes"], n)
        out = dict(n=np.uint32(n), pairs_sha=digest(b["pairs"]), nodes_sha=digest(nodes), unsorted_sha=digest(b["unsorted"]))
        if n <= 400:
            out["pairs"] = b["pairs"]; out["nodes"] = nodes; out["unsorted"] = b["unsorted"]
        hdr = np.frombuffer(scene[:96].tobytes(), dtype=L.SCENE_HEADER)[0]
        bb = hdr["modelsBoundingBox"]
        centre = (bb[0, :3] + bb[1, :3]) * 0.5
        ext = (bb[1, :3] - bb[0, :3])
        cams = [O.ref_demo_camera(scene, W, H),
                O.ref_camera_from_pose(W, H, 60.0, POSE_QUAT, (centre[0] + 0.9 * ext[2], centre[1] + 0.1 * ext[1], bb[0, 2] - 0.8 * ext[2]))]
        for k, cam in enumerate(cams):
            c = O.ref_trace_primary(cam, b["nodes"], n, scene)
            miss = c["nd"][:, 3] <= 0
            c["nd"][miss, :3] = 0            # uninitialised in the reference on a miss (BVH.h:353-354)
            out[f"cam{k}"] = cam
            out[f"contacts{k}"] = c
            out[f"colors{k}"] = O.ref_shade(cam, c, scene)
        np.savez_compressed(os.path.join(HERE, f"bvh_{name}.npz"), **out)
        # Two-Level Grid (densities 2.0 / 2.0, the reference's defaults): cell lists as digests (+ full arrays when small),
        # contacts for the same two cameras, all from the reference's own TwoLevelGrid.h functions.
        t = O.RefTlg(scene)
        e = t.export()
        tl = dict(sizes=np.array(t.sizes(), dtype=np.uint32), grid=e["grid"])
        for k in ("top_cells", "top_ranges", "top_pairs", "leaf_ranges", "leaf_pairs"):
            tl[k + "_sha"] = digest(e[k])
            if n <= 400:
                tl[k] = e[k]
        for k, cam in enumerate(cams):
            c = t.trace_primary(cam)
            miss = c["nd"][:, 3] <= 0
            c["nd"][miss] = 0
            tl[f"cam{k}"] = cam
            tl[f"contacts{k}"] = c
        np.savez_compressed(os.path.join(HERE, f"tlg_{name}.npz"), **tl)
        t.free()
        print(name, n, len(scene), [int((out[f'contacts{k}']['nd'][:, 3] > 0).sum()) for k in range(2)])


if __name__ == "__main__":
    main()
