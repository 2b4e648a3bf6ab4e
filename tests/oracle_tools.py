"""TEST INFRASTRUCTURE: Python access to oracle/ (the unmodified reference built into oracle/_ref, and the CPU
restatement oracle/_build/libcrt_oracle.so).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
reference arm may import this module; nothing in the product package does."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
HARNESS = os.path.join(REF_DIR, "ref_harness")
REF_CLI = os.path.join(REF_DIR, "crt_renderer_ref")
PORT_LIB = os.path.join(ORACLE_DIR, "_build", "libcrt_oracle.so")
REFERENCE = "/root/reference"
SCENES = os.path.join(REF_DIR, "scenes")


def have_reference_build() -> bool:
    return os.path.exists(HARNESS) and os.path.isdir(SCENES)


def ensure_reference_build() -> bool:
    """Build oracle/_ref when the reference sources are mounted (this container); on the GPU box the prebuilt
    binaries travel with the snapshot."""
    if have_reference_build():
        return True
    if os.path.isdir(os.path.join(REFERENCE, "src", "core")):
        subprocess.run(["make", "-j", "8", "ref"], cwd=ORACLE_DIR, check=True, stdout=subprocess.DEVNULL)
    return have_reference_build()


def ensure_port_build() -> bool:
    if not os.path.exists(PORT_LIB) or os.path.getmtime(PORT_LIB) < os.path.getmtime(os.path.join(ORACLE_DIR, "crt_oracle.cpp")):
        subprocess.run(["make", "port"], cwd=ORACLE_DIR, check=True, stdout=subprocess.DEVNULL)
    return os.path.exists(PORT_LIB)


def scene_file(relative: str) -> str:
    return os.path.join(SCENES, relative)


LOADABLE_SCENES = [
    "09-01-barycentric-coordinates/scene1.crtscene",
    "09-02-diffuse-smooth-shading/scene2.crtscene", "09-02-diffuse-smooth-shading/scene3.crtscene",
    "09-03-reflective/scene4.crtscene", "09-03-reflective/scene5.crtscene",
    *[f"11-01-refractive/scene{i}.crtscene" for i in range(9)],
    *[f"12-01-textures/scene{i}.crtscene" for i in range(5)],
    "13-01-optimizations/scene0.crtscene",
    "14-01-acceleration-tree/scene0.crtscene", "14-01-acceleration-tree/scene1.crtscene",
    "15-01-conclusion/scene0.crtscene", "15-01-conclusion/scene1.crtscene", "15-01-conclusion/scene2.crtscene",
]
UNLOADABLE_SCENES = [
    *[f"07-01-scene/scene{i}.crtscene" for i in range(5)], *[f"08-01-light/scene{i}.crtscene" for i in range(4)],
    "09-01-barycentric-coordinates/scene0.crtscene",
]


def scene_variant(relative: str, width=None, height=None, **settings) -> str:
    """JSON text of a bundled scene with image size / settings flags rewritten (never edits the reference)."""
    doc = json.load(open(scene_file(relative)))
    if width is not None:
        doc["settings"]["image_settings"]["width"] = int(width)
    if height is not None:
        doc["settings"]["image_settings"]["height"] = int(height)
    for k, v in settings.items():
        doc["settings"][k] = v
    return json.dumps(doc)


def run_harness(scene_text_or_path: str, asset_root: str | None = None, render=False, hits=False, rays=False, dump=False, stride=1,
                brute=False, max_depth=None, gi_rays=None, repeat=1):
    """Run oracle/_ref/ref_harness.  Returns dict with the JSON summary plus numpy arrays."""
    assert have_reference_build(), "oracle/_ref is not built"
    with tempfile.TemporaryDirectory() as tmp:
        if os.path.exists(scene_text_or_path):
            path = scene_text_or_path
        else:
            path = os.path.join(tmp, "scene.crtscene")
            with open(path, "w") as f:
                f.write(scene_text_or_path)
        cmd = [HARNESS, path]
        if asset_root:
            cmd += ["--asset-root", asset_root]
        files = {}
        for flag, on in (("render", render), ("hits", hits), ("rays", rays), ("dump", dump)):
            if on:
                files[flag] = os.path.join(tmp, flag + ".bin")
                cmd += ["--" + flag, files[flag]]
        if stride != 1:
            cmd += ["--stride", str(stride)]
        if brute:
            cmd += ["--brute"]
        if max_depth is not None:
            cmd += ["--max-depth", str(max_depth)]
        if gi_rays is not None:
            cmd += ["--gi-rays", str(gi_rays)]
        if repeat != 1:
            cmd += ["--repeat", str(repeat)]
        proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        out = {"returncode": proc.returncode}
        try:
            out.update(json.loads(proc.stdout.strip().splitlines()[-1]))
        except Exception:
            out["loaded"] = False
            out["stderr"] = proc.stderr
        if not out.get("loaded"):
            return out
        W, H = out["width"], out["height"]
        if render:
            out["rgb"] = np.fromfile(files["render"], np.float32).reshape(H, W, 3)
        if rays:
            sw, sh = (W + stride - 1) // stride, (H + stride - 1) // stride
            out["rays"] = np.fromfile(files["rays"], np.float32).reshape(sh, sw, 6)
        if hits:
            sw, sh = out["hits_w"], out["hits_h"]
            raw = np.fromfile(files["hits"], np.int32).reshape(4, sh, sw)
            out["tri"], out["t"], out["n_node"], out["n_tri"] = raw[0], raw[1].view(np.float32), raw[2], raw[3]
        if dump:
            out["dump"] = read_dump(files["dump"])
        return out


def read_dump(path: str) -> dict:
    """Parse ref_harness --dump (layout: oracle/ref_harness.cpp dump_scene)."""
    raw = open(path, "rb").read()
    off = 0

    def take(dtype, count):
        nonlocal off
        a = np.frombuffer(raw, dtype=dtype, count=count, offset=off)
        off += a.nbytes
        return a

    hdr = take(np.int32, 12)
    n_vert, n_nodes, n_refs, n_lights, n_tex, n_mat = (int(v) for v in hdr[:6])
    d = {"bucket_size": int(hdr[6]), "gi_on": int(hdr[7]), "reflections_on": int(hdr[8]), "refractions_on": int(hdr[9]),
         "width": int(hdr[10]), "height": int(hdr[11])}
    d["background"] = take(np.float32, 3)
    d["vertices"] = take(np.float32, n_vert * 9).reshape(n_vert, 9)
    node_dt = np.dtype([("min", np.float32, 3), ("max", np.float32, 3), ("c0", np.int32), ("c1", np.int32), ("parent", np.int32),
                        ("first", np.int32), ("count", np.int32)])
    d["nodes"] = take(node_dt, n_nodes)
    ref_dt = np.dtype([("v", np.int32, 3), ("material", np.int32), ("flags", np.int32), ("tri", np.int32), ("normal", np.float32, 3)])
    d["refs"] = take(ref_dt, n_refs)
    d["lights"] = take(np.float32, n_lights * 4).reshape(n_lights, 4)  # intensity, position
    mat_dt = np.dtype([("type", np.int32), ("texture", np.int32), ("ior", np.float32)])
    d["materials"] = take(mat_dt, n_mat)
    tex_dt = np.dtype([("type", np.int32), ("p", np.float32, 7), ("w", np.int32), ("h", np.int32)])
    d["textures"] = take(tex_dt, n_tex)
    d["texels"] = [take(np.float32, int(t["w"]) * int(t["h"]) * 3).reshape(-1, 3) for t in d["textures"] if t["type"] == 3]
    assert off == len(raw), (off, len(raw))
    return d


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def compare_flat_with_dump(flat, dump) -> list[str]:
    """Bit-exact comparison of OUR flattened scene (crt_b200.FlatScene) with the reference's loaded scene."""
    errs = []
    arr = flat.arrays()

    def same(name, ours, theirs):
        ours, theirs = np.asarray(ours), np.asarray(theirs)
        if ours.shape != theirs.shape:
            errs.append(f"{name}: shape {ours.shape} vs {theirs.shape}")
        elif ours.dtype.kind == "f":
            if not np.array_equal(bits(ours), bits(theirs)):
                errs.append(f"{name}: {int((bits(ours) != bits(theirs)).sum())} float words differ")
        elif not np.array_equal(ours, theirs):
            errs.append(f"{name}: {int((ours != theirs).sum())} entries differ")

    for k in ("bucket_size", "gi_on", "reflections_on", "refractions_on"):
        if int(getattr(flat, k)) != dump[k]:
            errs.append(f"{k}: {getattr(flat, k)} vs {dump[k]}")
    if (flat.camera.width, flat.camera.height) != (dump["width"], dump["height"]):
        errs.append("resolution differs")
    same("background", np.array(list(flat.background), np.float32), dump["background"])
    nv = flat.n_vertices
    if nv != len(dump["vertices"]):
        errs.append(f"vertex count {nv} vs {len(dump['vertices'])}")
        return errs
    same("vertex normals", arr["vert_normal"][:, :3], dump["vertices"][:, 3:6])
    same("vertex uv", arr["vert_uv"], dump["vertices"][:, 6:8])
    # tree
    nodes = dump["nodes"]
    if flat.n_nodes != len(nodes):
        errs.append(f"node count {flat.n_nodes} vs {len(nodes)}")
        return errs
    box = arr["box_nodes"]
    same("node min", box[:, 0:3].view(np.float32), nodes["min"])
    same("node max", box[:, 4:7].view(np.float32), nodes["max"])
    leaf = nodes["count"] > 0
    left = np.where(leaf, -nodes["first"] - 2, nodes["c0"])
    right = np.where(leaf, nodes["count"], nodes["c1"])
    same("node left", box[:, 3], left)
    same("node right", box[:, 7], right)
    kd = arr["kd_nodes"]
    same("kd leaf flag", kd[:, 3] == 3, leaf)
    same("kd leaf first", kd[leaf, 1], nodes["first"][leaf])
    same("kd leaf count", kd[leaf, 2], nodes["count"][leaf])
    same("kd child0", kd[~leaf, 1], nodes["c0"][~leaf])
    same("kd child1", kd[~leaf, 2], nodes["c1"][~leaf])
    # the split plane must equal the shared face of the two child boxes
    inner = np.nonzero(~leaf)[0]
    for child_col, face in (("c0", "max"), ("c1", "min")):
        has = nodes[child_col][inner] >= 0
        idx = inner[has]
        axis = kd[idx, 3]
        plane = nodes[face][nodes[child_col][idx], axis]
        same(f"kd split vs {child_col}.{face}", kd[idx, 0].view(np.float32), plane)
    # leaf references: triangle identity, flags, geometry
    refs = dump["refs"]
    if flat.n_refs != len(refs):
        errs.append(f"ref count {flat.n_refs} vs {len(refs)}")
        return errs
    our_tri = (arr["refs"] & 0x7FFFFFFF).astype(np.int64)
    same("ref triangle ids", our_tri, refs["tri"].astype(np.int64))
    same("ref cull bit", (arr["refs"] >> 31).astype(np.int32), (refs["flags"] >> 1) & 1)
    meta = arr["tri_meta"][our_tri]
    same("ref vertex indices", meta[:, :3], refs["v"])
    same("ref material", meta[:, 3] & 0xFFFFFF, refs["material"])
    same("ref smooth bit", (meta[:, 3] >> 24) & 1, refs["flags"] & 1)
    tv = arr["tri_verts"][our_tri]  # [n,3,4]
    same("ref face normals", tv[:, :, 3], refs["normal"])
    pos = dump["vertices"][:, 0:3]
    for k in range(3):
        same(f"ref v{k} positions", tv[:, k, :3], pos[refs["v"][:, k]])
    # lights / materials / textures
    same("light positions", arr["lights"][:, :3], dump["lights"][:, 1:4])
    same("light intensity", arr["lights"][:, 3], dump["lights"][:, 0])
    same("material type", arr["materials"][:, 0], dump["materials"]["type"])
    same("material texture", arr["materials"][:, 1], dump["materials"]["texture"])
    same("material ior", arr["materials"][:, 2].view(np.float32), dump["materials"]["ior"])
    tex = arr["textures"]
    dt = dump["textures"]
    same("texture type", tex[:, 0], dt["type"])
    texel_i = 0
    for i, t in enumerate(dt):
        ours = tex[i]
        c0, param, c1 = ours[4:7].view(np.float32), ours[7:8].view(np.float32), ours[8:11].view(np.float32)
        if t["type"] == 0:
            same(f"texture {i} albedo", c0, t["p"][:3])
        elif t["type"] in (1, 2):
            same(f"texture {i} colours", np.concatenate([c0, c1, param]), t["p"])
        else:
            if (int(ours[1]), int(ours[2])) != (int(t["w"]), int(t["h"])):
                errs.append(f"texture {i} size differs")
            else:
                n = int(t["w"]) * int(t["h"])
                same(f"texture {i} texels", arr["texels"][ours[3]: ours[3] + n, :3], dump["texels"][texel_i])
            texel_i += 1
    return errs


# ---------------------------------------------------------------------------------------------------------------
# the CPU restatement (oracle/crt_oracle.cpp) through ctypes.  Takes the same crt_flat_scene / crt_render_params
# structs as the product's C ABI, so a test hands BOTH sides literally the same bytes.
# ---------------------------------------------------------------------------------------------------------------
_port = None


def port_lib():
    global _port
    if _port is None:
        ensure_port_build()
        _port = C.CDLL(PORT_LIB)
        vp = C.c_void_p
        _port.crt_oracle_render.restype = C.c_int
        _port.crt_oracle_render.argtypes = [vp, vp, C.c_int, vp, vp, vp]
        _port.crt_oracle_primary_hits.restype = C.c_int
        _port.crt_oracle_primary_hits.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, vp]
        _port.crt_oracle_primary_rays.restype = C.c_int
        _port.crt_oracle_primary_rays.argtypes = [vp, vp]
    return _port


def _p(a):
    return C.c_void_p(a.ctypes.data)


def port_render(flat, params, threads: int = 0, primary: bool = True) -> dict:
    """Render with the port.  `flat` is a crt_b200.FlatScene, `params` a crt_b200.RenderParams (rank/world honoured)."""
    import sys
    crt = sys.modules["crt_b200"]
    lib = port_lib()
    W, H = flat.camera.width, flat.camera.height
    rows = len(crt.Renderer.owned_rows(H, params.rank, params.world, params.band_rows))
    rgb = np.zeros((rows, W, 3), np.float32)
    rgb8 = np.zeros((rows, W, 3), np.uint8)
    stats = crt.RenderStats()
    rc = lib.crt_oracle_render(C.byref(flat), C.byref(params), threads, _p(rgb), _p(rgb8), C.byref(stats))
    assert rc == 0
    out = {"rgb": rgb, "rgb8": rgb8, "stats": stats.as_dict()}
    if primary:
        out.update(port_primary_hits(flat, threads=threads))
    return out


def port_primary_hits(flat, brute: bool = False, threads: int = 0) -> dict:
    lib = port_lib()
    W, H = flat.camera.width, flat.camera.height
    tri = np.zeros((H, W), np.int32)
    t = np.zeros((H, W), np.float32)
    nn = np.zeros((H, W), np.int32)
    nt = np.zeros((H, W), np.int32)
    rc = lib.crt_oracle_primary_hits(C.byref(flat), int(brute), threads, _p(tri), _p(t), _p(nn), _p(nt))
    assert rc == 0
    return {"tri": tri, "t": t, "n_node": nn, "n_tri": nt}


def port_primary_rays(flat) -> np.ndarray:
    lib = port_lib()
    rays = np.zeros((flat.camera.height, flat.camera.width, 6), np.float32)
    assert lib.crt_oracle_primary_rays(C.byref(flat), _p(rays)) == 0
    return rays


# ---------------------------------------------------------------------------------------------------------------
# canonical byte strings of a loaded scene, computable from either side (for the committed digests)
# ---------------------------------------------------------------------------------------------------------------
def canonical_scene_bytes_from_dump(dump) -> dict:
    nodes, refs = dump["nodes"], dump["refs"]
    leaf = nodes["count"] > 0
    links = np.stack([np.where(leaf, -nodes["first"] - 2, nodes["c0"]), np.where(leaf, nodes["count"], nodes["c1"])], axis=1).astype(np.int32)
    return {
        "vertex_normals": np.ascontiguousarray(dump["vertices"][:, 3:6]).tobytes(),
        "vertex_uvs": np.ascontiguousarray(dump["vertices"][:, 6:8]).tobytes(),
        "node_boxes": np.ascontiguousarray(np.concatenate([nodes["min"], nodes["max"]], axis=1)).tobytes(),
        "node_links": links.tobytes(),
        "ref_triangles": np.ascontiguousarray(refs["tri"].astype(np.int32)).tobytes(),
        "ref_vertices": np.ascontiguousarray(refs["v"].astype(np.int32)).tobytes(),
        "ref_normals": np.ascontiguousarray(refs["normal"]).tobytes(),
        "lights": np.ascontiguousarray(dump["lights"]).tobytes(),
    }


def canonical_scene_bytes_from_flat(flat) -> dict:
    a = flat.arrays()
    box = a["box_nodes"]
    tri = (a["refs"] & 0x7FFFFFFF).astype(np.int64)
    return {
        "vertex_normals": np.ascontiguousarray(a["vert_normal"][:, :3]).tobytes(),
        "vertex_uvs": np.ascontiguousarray(a["vert_uv"]).tobytes(),
        "node_boxes": np.ascontiguousarray(np.concatenate([box[:, 0:3], box[:, 4:7]], axis=1)).tobytes(),
        "node_links": np.ascontiguousarray(np.stack([box[:, 3], box[:, 7]], axis=1).astype(np.int32)).tobytes(),
        "ref_triangles": tri.astype(np.int32).tobytes(),
        "ref_vertices": np.ascontiguousarray(a["tri_meta"][tri][:, :3].astype(np.int32)).tobytes(),
        "ref_normals": np.ascontiguousarray(a["tri_verts"][tri][:, :, 3]).tobytes(),
        "lights": np.ascontiguousarray(np.concatenate([a["lights"][:, 3:4], a["lights"][:, :3]], axis=1)).tobytes(),
    }


def golden_path(rel: str) -> str:
    return os.path.join(ROOT, "tests", "golden", rel.replace("/", "__").replace(".crtscene", "") + ".npz")


def golden_size(rel: str):
    return (48, 48) if rel.startswith("15-01-conclusion/scene2") else (96, 54)
