"""Generates tests/golden/obj/*.npz: what the reference's loader chain (its vendored tinyobjloader,
flattened like load_meshes, src/main.cpp:143-223) produces for the hand-written fixtures.
Needs oracle/_ref/libtinyobj_ref.so, i.e. a container where /root/reference exists."""
import glob, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import oracle_py as O

for p in sorted(glob.glob(os.path.join(HERE, "obj", "*.obj"))):
    r = O.ref_load_obj(p)
    np.savez_compressed(p[:-4] + ".npz", **r)
    print(os.path.basename(p), {k: v.shape for k, v in r.items()})
