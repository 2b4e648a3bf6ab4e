"""Host logic: the .crtscene loader, the tree builder and the flattener against the reference (live when oracle/_ref is
built, and against the committed digests generated from it)."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle_tools as otools

DIGESTS = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "scene_digests.json")))
needs_scenes = pytest.mark.skipif(not (otools.have_reference_build() or os.path.isdir("/root/reference/src")),
                                  reason="scene fixtures live in oracle/_ref/scenes (built where /root/reference is mounted)")


@needs_scenes
@pytest.mark.parametrize("rel", otools.LOADABLE_SCENES)
def test_flattened_scene_matches_reference_digests(crt, ot, rel):
    flat_owner = crt.Scene.load_file(ot.scene_file(rel))
    flat = flat_owner.flatten()
    got = {k: hashlib.sha256(v).hexdigest() for k, v in ot.canonical_scene_bytes_from_flat(flat).items()}
    assert got == DIGESTS[rel]


@needs_scenes
@pytest.mark.parametrize("rel", ["14-01-acceleration-tree/scene1.crtscene", "15-01-conclusion/scene2.crtscene", "12-01-textures/scene4.crtscene",
                                 "11-01-refractive/scene8.crtscene", "09-03-reflective/scene4.crtscene"])
def test_flattened_scene_matches_live_reference_dump(crt, ot, rel):
    if not ot.have_reference_build():
        pytest.skip("oracle/_ref not built")
    scene = crt.Scene.load_file(ot.scene_file(rel))
    ref = ot.run_harness(ot.scene_file(rel), dump=True)
    assert ref["loaded"]
    assert ot.compare_flat_with_dump(scene.flatten(), ref["dump"]) == []


@needs_scenes
@pytest.mark.parametrize("rel", otools.UNLOADABLE_SCENES)
def test_scenes_the_reference_rejects_are_rejected(crt, ot, rel):
    # missing materials / lights / material_index
    with pytest.raises(crt.CrtError):
        crt.Scene.load_file(ot.scene_file(rel))


MINIMAL = {
    "settings": {"background_color": [0, 0.5, 0], "image_settings": {"width": 8, "height": 4}},
    "camera": {"matrix": [1, 0, 0, 0, 1, 0, 0, 0, 1], "position": [0, 0, 0]},
    "lights": [{"intensity": 10, "position": [0, 1, 0]}],
    "materials": [{"type": "diffuse", "albedo": [1, 0, 0], "smooth_shading": False}],
    "objects": [{"material_index": 0, "vertices": [-1, -1, -3, 1, -1, -3, 0, 1, -3], "triangles": [0, 1, 2]}],
}


def _load(crt, doc):
    return crt.Scene.load_json(json.dumps(doc) if not isinstance(doc, str) else doc)


def test_minimal_scene_and_defaults(crt):
    flat_owner = _load(crt, MINIMAL)
    flat = flat_owner.flatten()
    assert (flat.n_triangles, flat.n_vertices, flat.n_nodes, flat.n_refs, flat.n_lights) == (1, 3, 1, 1, 1)
    assert (flat.gi_on, flat.reflections_on, flat.refractions_on, flat.bucket_size) == (0, 1, 1, 24)  # crt_json.cpp:616, crt_scene.h:16
    assert flat.camera.tan_half_fov == pytest.approx(1.0, rel=1e-6)  # fov default 90 degrees
    arr = flat.arrays()
    assert np.allclose(arr["tri_verts"][0, :, 3], [0, 0, 1])  # face normal of a CCW triangle facing +z
    assert arr["materials"][0, 1] == 0 and flat.n_textures == 1  # inline colour -> implicit albedo texture (crt_json.cpp:507-515)


@pytest.mark.parametrize("mutate", [
    lambda d: d.pop("materials"), lambda d: d.pop("lights"), lambda d: d.pop("objects"), lambda d: d.pop("camera"),
    lambda d: d.pop("settings"), lambda d: d["settings"].pop("image_settings"), lambda d: d["settings"].pop("background_color"),
    lambda d: d.update(materials=[]), lambda d: d["objects"][0].pop("material_index"),
    lambda d: d["objects"][0].update(material_index=5), lambda d: d["objects"][0].update(triangles=[0, 1]),
    lambda d: d["objects"][0].update(triangles=[0, 1, 2.5]), lambda d: d["objects"][0].update(triangles=[0, 1, 7]),
    lambda d: d["objects"][0].update(vertices=[0, 1]), lambda d: d["objects"][0].update(uvs=[0, 0, 0]),
    lambda d: d["materials"][0].pop("smooth_shading"), lambda d: d["materials"][0].update(type="glossy"),
    lambda d: d["materials"][0].update(albedo="no such texture"), lambda d: d["camera"].update(matrix=[1, 0, 0]),
    lambda d: d["settings"]["image_settings"].update(width=8.5), lambda d: d["settings"].update(gi_on=1),
    lambda d: d["camera"].update(fov_degrees="wide"), lambda d: d["lights"][0].pop("intensity"),
])
def test_validation_failures_are_parse_errors(crt, mutate):
    doc = json.loads(json.dumps(MINIMAL))
    mutate(doc)
    with pytest.raises(crt.CrtError, match="-4"):
        _load(crt, doc)


@pytest.mark.parametrize("text", ["", "   ", "{", "[1,2", '{"a":1,}', '{"a":1} x', "{'a':1}", '{"a":01}', '{"a":1.}', '{"a":.5}', '{"a":1e}',
                                  '{"a":NaN}', '{"a":1e999}', '{"a":"\\x"}', '{"a":"\\ud800"}', "[1 2]", "nul"])
def test_malformed_json_is_rejected(crt, text):
    with pytest.raises(crt.CrtError):
        crt.Scene.load_json(text)


def test_number_conversion_follows_rapidjson(crt):
    # every number goes through double -> float exactly as rapidjson's GetFloat(); spot values incl. long mantissas, exponents, -0
    values = ["0.1", "-0.30000001192092896", "1e-3", "2.5E+2", "123456789.123456789", "0.000001", "16777217", "-0", "-0.0",
              "3.4028234663852886e38", "1.17549435e-38", "0.333333333333333314829616256247", "4294967296", "18446744073709551616"]
    doc = json.loads(json.dumps(MINIMAL))
    text = json.dumps(doc).replace('"position": [0, 0, 0]', '"position": [0, 0, 0], "fov_degrees": 60')
    for v in values:
        t = text.replace('"intensity": 10', f'"intensity": {v}')
        flat_owner = crt.Scene.load_json(t)
        got = flat_owner.flatten().arrays()["lights"][0, 3]
        want = np.float32(float(v))
        assert got.tobytes() == want.tobytes() or (v == "-0" and got == 0 and not np.signbit(got)), (v, got, want)


def test_texture_list_failure_voids_all_textures(crt):
    # crt_json.cpp:582-588: a malformed entry drops the WHOLE list; inline-colour materials still load
    doc = json.loads(json.dumps(MINIMAL))
    doc["textures"] = [{"name": "ok", "type": "albedo", "albedo": [1, 1, 1]}, {"name": "bad", "type": "edges"}]
    flat_owner = _load(crt, doc)
    assert flat_owner.flatten().n_textures == 1
    doc["materials"][0]["albedo"] = "ok"
    with pytest.raises(crt.CrtError):
        _load(crt, doc)


def test_vertex_normals_are_renormalised_after_every_object(crt):
    # crt_mesh.cpp:27-29: unreferenced vertices become NaN
    doc = json.loads(json.dumps(MINIMAL))
    doc["objects"][0]["vertices"] += [5, 5, 5]
    n = _load(crt, doc).flatten().arrays()["vert_normal"]
    assert np.isnan(n[3, :3]).all() and np.allclose(n[:3, :3], [0, 0, 1])


def test_tree_rules_on_a_synthetic_grid(crt):
    # leaf rule (<= 16 triangles or depth > 39), straddlers duplicated, absent child = -1, DFS pre-order numbering
    rng = np.random.default_rng(7)
    pts = rng.uniform(-4, 4, size=(300, 3)).astype(np.float32)
    verts, tris = [], []
    for i, p in enumerate(pts):
        verts += [*p, *(p + [0.3, 0, 0]), *(p + [0, 0.3, 0])]
        tris += [3 * i, 3 * i + 1, 3 * i + 2]
    doc = json.loads(json.dumps(MINIMAL))
    doc["objects"][0].update(vertices=[float(v) for v in verts], triangles=tris)
    flat = _load(crt, doc).flatten()
    a = flat.arrays()
    box = a["box_nodes"]
    left, right = box[:, 3], box[:, 7]
    leaf = left <= -2
    assert leaf.any() and (~leaf).any()
    assert (right[leaf] > 0).all()
    assert flat.n_refs == int(right[leaf].sum()) and flat.n_refs >= flat.n_triangles
    first = -left[leaf] - 2
    assert (np.diff(first) > 0).all() and first[0] == 0  # leaves own consecutive slices in DFS order
    for i in np.nonzero(~leaf)[0]:
        kids = [k for k in (left[i], right[i]) if k != -1]
        assert kids and all(k > i for k in kids)
        if left[i] != -1:
            assert left[i] == i + 1  # child0 directly follows its parent (pre-order)
    # every triangle is referenced, and only triangles whose box overlaps the leaf box
    refs = a["refs"] & 0x7FFFFFFF
    assert set(refs.tolist()) == set(range(flat.n_triangles))
    mins, maxs = box[:, 0:3].view(np.float32), box[:, 4:7].view(np.float32)
    tv = a["tri_verts"][:, :, :3]
    for i in np.nonzero(leaf)[0][:50]:
        for t in refs[-left[i] - 2: -left[i] - 2 + right[i]]:
            assert (tv[t].min(0) <= maxs[i]).all() and (tv[t].max(0) >= mins[i]).all()


@pytest.mark.parametrize("rel", ["14-01-acceleration-tree/scene1.crtscene", "15-01-conclusion/scene2.crtscene", "11-01-refractive/scene3.crtscene"])
def test_walk_tree_is_the_reference_tree_with_subtrees_merged(crt, ot, rel, monkeypatch):
    """CRT_WALK_TREE=reference: fast_nodes / fast_refs have the split planes and topology of the reference tree,
    every merged leaf holds exactly the triangles of the reference subtree it replaces."""
    monkeypatch.setenv("CRT_WALK_TREE", "reference")
    scene = crt.Scene.load_file(ot.scene_file(rel))
    flat = scene.flatten()
    a = flat.arrays()
    kd, refs, fast, frefs = a["kd_nodes"], a["refs"] & 0x7FFFFFFF, a["fast_nodes"], a["fast_refs"]
    assert flat.n_fast_nodes == len(fast) <= flat.n_nodes and flat.fast_max_depth <= flat.max_depth

    def subtree_triangles(i):
        if kd[i, 3] == 3:
            return set(refs[kd[i, 1]:kd[i, 1] + kd[i, 2]].tolist())
        out = set()
        for c in (kd[i, 1], kd[i, 2]):
            if c >= 0:
                out |= subtree_triangles(c)
        return out

    # walk both trees in lock step from the roots
    todo, seen_fast = [(0, 0)], 0
    while todo:
        i, j = todo.pop()
        seen_fast += 1
        if fast[j, 3] == 3:
            lst = (frefs[fast[j, 1]:fast[j, 1] + fast[j, 2]] & 0x7FFFFFFF).tolist()
            assert lst == sorted(set(lst)), "leaf lists are ascending and duplicate-free"
            assert set(lst) == subtree_triangles(i)
            cull = frefs[fast[j, 1]:fast[j, 1] + fast[j, 2]] >> 31
            assert np.array_equal(cull, (a["tri_meta"][lst, 3] >> 25) & 1)
        else:
            assert kd[i, 3] == fast[j, 3] and kd[i, 0] == fast[j, 0], "same axis and split plane"
            for col in (1, 2):
                assert (kd[i, col] >= 0) == (fast[j, col] >= 0)
                if kd[i, col] >= 0:
                    todo.append((kd[i, col], fast[j, col]))
            if fast[j, 1] >= 0:
                assert fast[j, 1] == j + 1, "child0 follows its parent (DFS pre-order)"
    assert seen_fast == len(fast)


@needs_scenes
@pytest.mark.parametrize("rel", ["14-01-acceleration-tree/scene1.crtscene", "15-01-conclusion/scene2.crtscene", "11-01-refractive/scene3.crtscene",
                                 "12-01-textures/scene4.crtscene", "14-01-acceleration-tree/scene0.crtscene"])
def test_sah_walk_tree_invariants(crt, ot, rel):
    """The default walk tree (core/crt_kdbuild.cpp): a k-d tree over the root box whose leaves list every triangle
    that reaches into their cell -- the property the closest-hit walk needs, whatever the split planes are."""
    scene = crt.Scene.load_file(ot.scene_file(rel))
    flat = scene.flatten()
    a = flat.arrays()
    fast, frefs = a["fast_nodes"], a["fast_refs"]
    tv = a["tri_verts"][:, :, :3]
    lo, hi = tv.min(1), tv.max(1)
    assert flat.n_fast_nodes == len(fast) and flat.n_fast_refs == len(frefs) and flat.fast_max_depth + 2 <= 48
    root_lo, root_hi = np.array(list(flat.root_min), dtype=np.float32), np.array(list(flat.root_max), dtype=np.float32)
    seen_nodes, listed, max_depth = 0, set(), 0
    todo = [(0, root_lo, root_hi, np.arange(flat.n_triangles), 0)]
    while todo:
        j, clo, chi, members, depth = todo.pop()
        seen_nodes += 1
        max_depth = max(max_depth, depth)
        split, c0, c1, axis = fast[j, 0].view(np.float32), int(fast[j, 1]), int(fast[j, 2]), int(fast[j, 3])
        if axis == 3:
            lst = frefs[c0:c0 + c1]
            ids = (lst & 0x7FFFFFFF).astype(np.int64)
            assert c1 > 0 and np.array_equal(ids, members), "leaf list = the triangles routed here, ascending"
            assert np.array_equal(lst >> 31, (a["tri_meta"][ids, 3] >> 25) & 1)
            # geometric reading of the same rule: whatever overlaps the cell's interior is listed
            inside = np.nonzero(((lo < chi) & (hi > clo)).all(1))[0]
            assert set(inside.tolist()) <= set(ids.tolist())
            listed |= set(ids.tolist())
            continue
        assert axis in (0, 1, 2) and clo[axis] < split < chi[axis] or flat.n_triangles == 0
        flat_in_plane = (lo[members, axis] == split) & (hi[members, axis] == split)
        left = members[(lo[members, axis] < split) | flat_in_plane]
        right = members[(hi[members, axis] > split) | flat_in_plane]
        assert (c0 >= 0) == (len(left) > 0) and (c1 >= 0) == (len(right) > 0), "absent child <=> nothing on that side"
        if c0 >= 0:
            assert c0 == j + 1, "child0 follows its parent (DFS pre-order)"
            h = chi.copy(); h[axis] = split
            todo.append((c0, clo, h, left, depth + 1))
        if c1 >= 0:
            l = clo.copy(); l[axis] = split
            todo.append((c1, l, chi, right, depth + 1))
    assert seen_nodes == len(fast) and max_depth == flat.fast_max_depth
    assert listed == set(range(flat.n_triangles)), "every triangle is reachable"


@pytest.mark.parametrize("rel", ["14-01-acceleration-tree/scene1.crtscene", "11-01-refractive/scene3.crtscene"])
def test_flatten_without_the_reference_tree(crt, ot, rel):
    """crt_scene_flatten_ex without CRT_FLATTEN_REFERENCE_TREE: no reference tree (it is not even built), everything
    else -- root box, triangles, walk tree -- identical; asking for the tree later fills it in and keeps the first view
    valid"""
    with open(ot.scene_file(rel)) as f:
        text = f.read()
    root = os.path.dirname(ot.scene_file(rel))
    lean_scene, full_scene = crt.Scene.load_json(text, root), crt.Scene.load_json(text, root)
    lean, full = lean_scene.flatten(reference_tree=False), full_scene.flatten()
    assert lean.n_nodes == 0 and lean.n_refs == 0 and lean.max_depth == 0 and full.n_nodes > 0
    assert list(lean.root_min) == list(full.root_min) and list(lean.root_max) == list(full.root_max)
    a, b = lean.arrays(), full.arrays()
    for key in ("tri_verts", "tri_meta", "vert_normal", "vert_uv", "fast_nodes", "fast_refs", "lights", "materials", "textures", "texels"):
        assert np.array_equal(otools.bits(a[key]) if a[key].dtype == np.float32 else a[key], otools.bits(b[key]) if b[key].dtype == np.float32 else b[key]), key
    later = lean_scene.flatten()  # upgrade in place
    c = later.arrays()
    for key in ("box_nodes", "kd_nodes", "refs"):
        assert np.array_equal(c[key], b[key]), key
    assert later.n_nodes == full.n_nodes and later.max_depth == full.max_depth
    assert np.array_equal(lean.arrays()["fast_nodes"], b["fast_nodes"])  # the first view still reads the same arrays


def _pack(crt, nodes):
    nodes = np.ascontiguousarray(nodes, dtype=np.int32)
    out = np.zeros(len(nodes), dtype=np.uint64)
    rc = crt.lib.crt_cuda_pack_walk_tree(nodes.ctypes.data, len(nodes), out.ctypes.data)
    return rc, out


@needs_scenes
@pytest.mark.parametrize("walk_tree", ["sah", "reference"])
@pytest.mark.parametrize("rel", ["14-01-acceleration-tree/scene1.crtscene", "15-01-conclusion/scene2.crtscene", "14-01-acceleration-tree/scene0.crtscene",
                                 "12-01-textures/scene4.crtscene"])
def test_device_node_records_hold_the_same_tree(crt, ot, rel, walk_tree, monkeypatch):
    """crt_cuda_pack_walk_tree (what crt_cuda_upload_scene puts in device memory): 8 bytes per node, children side by
    side.  Walked together with the 16-byte tree it came from, it has the same planes, axes, absent children and leaf
    lists; every node is used exactly once; the root is node 0."""
    if walk_tree == "reference":
        monkeypatch.setenv("CRT_WALK_TREE", "reference")
    flat = crt.Scene.load_file(ot.scene_file(rel)).flatten(reference_tree=walk_tree == "reference")
    fast = flat.arrays()["fast_nodes"]
    rc, packed = _pack(crt, fast)
    assert rc == 0
    lo, hi = (packed & 0xFFFFFFFF).astype(np.uint32), (packed >> 32).astype(np.uint32)
    used = np.zeros(len(fast), dtype=np.int32)
    todo = [(0, 0)]
    while todo:
        old, new = todo.pop()
        used[new] += 1
        split_bits, a, b, axis = (int(v) for v in fast[old])
        assert int(hi[new] & 3) == axis
        if axis == 3:
            assert int(lo[new]) == a and int(hi[new] >> 2) == b, "leaf: first reference and count"
            continue
        assert int(lo[new]) == split_bits & 0xFFFFFFFF, "split plane bit for bit"
        has_a, has_b, child = int(hi[new] >> 2) & 1, int(hi[new] >> 3) & 1, int(hi[new] >> 4)
        assert has_a == (a >= 0) and has_b == (b >= 0)
        assert not (has_a or has_b) or 0 < child and child + has_a + has_b <= len(fast)
        if has_a:
            todo.append((a, child))
        if has_b:
            todo.append((b, child + has_a))
    assert (used == 1).all(), "a permutation of the nodes"


def test_device_node_packing_refuses_broken_trees(crt):
    leaf = [0, 0, 1, 3]
    assert _pack(crt, [[0, 1, 2, 0], leaf, leaf])[0] == 0
    assert _pack(crt, [[0, 1, 1, 0], leaf])[0] == crt.ERR_INVALID  # a node reachable twice: does not unfold into n records
    assert _pack(crt, [[0, 1, 2, 5], leaf, leaf])[0] == crt.ERR_INVALID  # axis out of range
    assert _pack(crt, [[0, 1, 7, 1], leaf, leaf])[0] == crt.ERR_INVALID  # child index out of range
    assert _pack(crt, [[0, 0, -1, 3]])[0] == crt.ERR_INVALID            # leaf with a negative count
    assert crt.lib.crt_cuda_pack_walk_tree(None, 1, None) == crt.ERR_INVALID
    rc, one = _pack(crt, [[7, -1, 2, 1], leaf, leaf])  # child a absent: b takes the first slot
    assert rc == 0 and int(one[0] >> 32) == (1 | 0 << 2 | 1 << 3 | 1 << 4)
