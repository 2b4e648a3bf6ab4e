"""Generates tests/golden/*.npz and scene_digests.json FROM THE UNMODIFIED REFERENCE (oracle/_ref/ref_harness, built
from /root/reference by oracle/Makefile).  Run in the build container:  python tests/golden/make_golden.py
Each npz holds, for one bundled scene re-sized to 96x54 (15-01/scene2: 48x48, it traces 73 rays per pixel):
  rgb   float32 [H,W,3]  crt::render_image output          tri  int32 [H,W] parse-order id of the primary hit (-1 miss)
  t     float32 [H,W]    primary hit distance               n_node / n_tri int32 [H,W]  box / triangle tests per primary ray
  rays  float32 [H,W,6]  Camera::generate_ray (origin, direction)
scene_digests.json holds sha256 digests of the reference's loaded scene (vertex normals/uvs, tree boxes and links,
leaf references) at native size, in the canonical byte order of oracle_tools.canonical_scene_bytes."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_tools as ot  # noqa: E402


def golden_size(rel):
    return (48, 48) if rel.startswith("15-01-conclusion/scene2") else (96, 54)


def main():
    assert ot.ensure_reference_build(), "needs /root/reference"
    digests = {}
    for rel in ot.LOADABLE_SCENES:
        w, h = golden_size(rel)
        root = os.path.dirname(ot.scene_file(rel))
        out = ot.run_harness(ot.scene_variant(rel, w, h), asset_root=root, render=True, hits=True, rays=True)
        name = rel.replace("/", "__").replace(".crtscene", "")
        np.savez_compressed(os.path.join(HERE, name + ".npz"), rgb=out["rgb"], tri=out["tri"], t=out["t"], n_node=out["n_node"],
                            n_tri=out["n_tri"], rays=out["rays"])
        dump = ot.run_harness(ot.scene_file(rel), dump=True)["dump"]
        digests[rel] = {k: hashlib.sha256(v).hexdigest() for k, v in ot.canonical_scene_bytes_from_dump(dump).items()}
        print(rel, "ok")
    json.dump(digests, open(os.path.join(HERE, "scene_digests.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
