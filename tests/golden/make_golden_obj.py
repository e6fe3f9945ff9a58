"""Packs the hand-written OBJ fixtures under tests/golden/obj/ with the REFERENCE's own loader (its vendored tinyobj +
the restated Scene::load packing loop, oracle/_ref/libclrt_ref.so) and commits the bytes: tests/golden/obj/packed.npz.
Run in the build container only (needs /root/reference):  python tests/golden/make_golden_obj.py"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle_py as O  # noqa: E402

CASES = {                       # name -> (lights, obj files)
    "quirks": ([[0.0, 10.0, -330.0, 1000.0]], ["quirks.obj"]),
    "nomtl": ([[1.5, 2.5, -3.5, 40.0], [0.0, 0.0, 0.0, 1.0]], ["nomtl.obj"]),
    "two_models": ([[0.0, 10.0, -330.0, 1000.0]], ["quirks.obj", "second.obj"]),
    "no_lights": ([], ["second.obj"]),
}


def main():
    O.build(("ref",))
    out = {}
    for name, (lights, objs) in CASES.items():
        out[name] = O.ref_scene_pack(np.array(lights, dtype=np.float32).reshape(-1, 4), [os.path.join(HERE, "obj", o) for o in objs])
        print(name, len(out[name]), "bytes")
    np.savez_compressed(os.path.join(HERE, "obj", "packed.npz"), **out)


if __name__ == "__main__":
    main()
