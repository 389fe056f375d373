"""Generates the committed golden fixtures from the REAL reference intersector (Embree 3.6.1 built
from /root/reference by oracle/ref_embree.mk).  Run in the authoring container only:
    python tests/golden/make_golden.py
The fixtures travel to the GPU box; /root/reference and (possibly) oracle/_ref do not have to."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "smallpt-enoki-optix_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import oracle_py as O  # noqa: E402
import scenes  # noqa: E402
import util  # noqa: E402


def main():
    assert O.ref_available(), "build oracle/_ref first (make -C oracle ref)"
    # 1. hit vectors: Embree on seeded rays, two small scenes
    for name, sd in (("cornell8", scenes.cornell(8)), ("knob40", scenes.knob(40, 16))):
        om = util.oracle_mesh(sd)
        o1, d1 = util.random_rays(sd, 3000, seed=101, inside=True)
        o2, d2 = util.random_rays(sd, 1000, seed=102, inside=False)
        o, d = np.concatenate([o1, o2]), np.concatenate([d1, d2])
        hid, t, u, v = om.intersect("embree", o, d)
        np.savez_compressed(os.path.join(HERE, f"hits_{name}.npz"), o=o, d=d, id=hid, t=t, u=u, v=v)
        print(name, "hit rate", (hid >= 0).mean())
    # 2. a tiny image through the restated loop with the reference's RNG (PCG32) on Embree
    sd = scenes.cornell(8, colored=True)
    om, cam = util.oracle_mesh(sd), util.oracle_camera(sd)
    film, seg = om.render("embree", cam, 32, 32, 0, 16, 5, rng_mode=O.RNG_PCG32)
    np.savez_compressed(os.path.join(HERE, "image_cornell8_pcg32.npz"), film=film, segments=seg)
    film2, seg2 = om.render("embree", cam, 32, 32, 0, 16, 5, rng_mode=O.RNG_PHILOX, seed=2024)
    np.savez_compressed(os.path.join(HERE, "image_cornell8_philox.npz"), film=film2, segments=seg2)
    print("images", film.mean() / 16, seg, film2.mean() / 16, seg2)


if __name__ == "__main__":
    main()
