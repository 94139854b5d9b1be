"""Synthetic shape specs and polygonizer settings for the BASELINE configs.

Pure data generators: they build the two JSON strings the geometry API takes.  Matrices are row-major 4x4
(`[R|t]` in the first 12 entries) exactly as the reference's mp5 documents carry them
(/root/reference/js_iteration_2/examples/js/example_objects.js:39).
"""
import json
import numpy as np

EYE = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]


def _mat(R, t):
    m = np.eye(4)
    m[:3, :3] = R
    m[:3, 3] = t
    # round-trip through float32 text so every consumer parses the same binary32 values
    return [float(np.float32(v)) for v in m.reshape(-1)]


def _rot_axis(axis, deg):
    axis = np.asarray(axis, dtype=np.float64)
    axis = axis / np.linalg.norm(axis)
    a = np.deg2rad(deg)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * (K @ K)


def box(h):
    return {"xmin": -h, "xmax": h, "ymin": -h, "ymax": h, "zmin": -h, "zmax": h}


def mc_settings(resolution, half, iters=0, c=1.0, projection=0, qem=0, repeats=1, **gpu):
    s = {"resolution": int(resolution), "box": box(half), "ignore_root_matrix": 0, "subdiv": {"enabled": 0},
         "overall_repeats": int(repeats), "vresampl": {"iters": int(iters), "c": float(c)},
         "projection": {"enabled": int(projection)}, "qem": {"enabled": int(qem)}}
    if gpu:
        s["gpu"] = gpu
    return json.dumps(s)


def c1_sphere():
    """C1: sphere r=0.5 requested as an ellipsoid."""
    return json.dumps({"type": "ellipsoid", "matrix": EYE})


def tiny_sphere(scale=0.03, center=(0.0, 0.0, 0.0)):
    """A sphere of radius scale/2: polygonized in a box of a few hundredths its mean edge falls below 0.004 (the projection's
    alpha list has fewer than the ten entries direction types 1-6 ask for) or below 0.001 (the list is empty and no centroid
    finds a candidate) — make_alpha_list, centroids_projection.cpp:144-194."""
    return json.dumps({"type": "ellipsoid", "matrix": [scale, 0, 0, center[0], 0, scale, 0, center[1], 0, 0, scale, center[2], 0, 0, 0, 1]})


def c2_cube_minus_cylinder():
    """C2: Difference[cube, cylinder(diag(0.6,0.6,1.4))] under Rot((1,1,0),30deg)*0.9 + t."""
    R = _rot_axis([1, 1, 0], 30.0) * 0.9
    node = _mat(R, [0.05, -0.03, 0.02])
    cyl = _mat(np.diag([0.6, 0.6, 1.4]), [0, 0, 0])
    return json.dumps({"type": "Difference", "matrix": node,
                       "children": [{"type": "cube", "matrix": EYE}, {"type": "cylinder", "matrix": cyl}]})


def c3_metaballs():
    """C3(i): Union of 16 meta_balls nodes (4 blobs each) on a 4x2x2 lattice, spacing 0.4."""
    kids = []
    i = 0
    for ix in range(4):
        for iy in range(2):
            for iz in range(2):
                t = [(ix - 1.5) * 0.4, (iy - 0.5) * 0.4, (iz - 0.5) * 0.4]
                kids.append({"type": "meta_balls", "time": float(np.float32(0.1 + 0.37 * i)), "matrix": _mat(np.eye(3), t)})
                i += 1
    return json.dumps({"type": "Union", "matrix": EYE, "children": kids})


def c3_sum_blend(n_balls=64, strength=0.2, subtract=0.75, seed=64):
    """C3(ii), the perf variant of SURVEY §8(d): ONE meta_balls node whose 64 blobs are summed (the "balls" extension key — the
    reference's factory always builds four blobs from `time`, so this spec is not expressible there; the oracle evaluates it with
    the reference's own loop).  Centres default_rng(64).uniform(-0.5, 0.5); the reference removed the val > 0 clamp, so every ball
    contributes -subtract/100 far away and `subtract` is scaled with the count (12 * 4 / 64 = 0.75); strength 0.2 gives an iso-surface
    of area 5.2 inside [-0.63, 0.63]^3 (SURVEY asks for 5-8)."""
    rng = np.random.default_rng(seed)
    cen = rng.uniform(-0.5, 0.5, size=(n_balls, 3))
    flat = []
    for c in cen:
        flat += [float(np.float32(c[0])), float(np.float32(c[1])), float(np.float32(c[2])), float(np.float32(strength)), float(np.float32(subtract))]
    return json.dumps({"type": "meta_balls", "time": 0.1, "matrix": EYE, "balls": flat})


def c4_deep_chain(n_prims=32, seed=1234):
    """C4: left-deep chain of n_prims transformed primitives, ops cycle [Union x3, Difference x1]."""
    rng = np.random.default_rng(seed)
    types = ["ellipsoid", "cylinder", "cone", "ellipsoid"]
    ops = ["Union", "Union", "Union", "Difference"]

    def prim(k):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        R = q * rng.uniform(0.15, 0.35)
        t = rng.uniform(-0.7, 0.7, size=3)
        return {"type": types[k % 4], "matrix": _mat(R, t)}

    def small_affine():
        axis = rng.normal(size=3)
        R = _rot_axis(axis, rng.uniform(-10, 10)) * rng.uniform(0.95, 1.05)
        return _mat(R, [0, 0, 0])

    node = prim(0)
    for k in range(1, n_prims):
        node = {"type": ops[(k - 1) % 4], "matrix": small_affine(), "children": [node, prim(k)]}
    return json.dumps(node)


def demo_moon():
    """MOON demo (nested Difference of cylinders), examples/js/example_objects.js:22-70; the root matrix is scaled by 1/60 so
    that the 42.5 x 49.6 x 10 plate fits a unit box (children's matrices are relative to their parents and stay as they are)."""
    def sc(m, scale=1.0):
        m = np.array(m, dtype=np.float64).reshape(4, 4)
        m[:3, :] *= scale
        return [float(np.float32(v)) for v in m.reshape(-1)]
    inner = {"type": "Difference", "matrix": sc([2.38, 0, 0, 0.36, 0, 2.38, 0, 0.56, 0, 0, 2.38, 6.9, 0, 0, 0, 1]),
             "children": [{"type": "cylinder", "matrix": sc([10, 0, 0, 0, 0, 10, 0, 0, 0, 0, 10, 0, 0, 0, 0, 1])},
                          {"type": "cylinder", "matrix": sc([10, 0, 0, 0.66, 0, 10, 0, 6.22, 0, 0, 10, 0.00001, 0, 0, 0, 1])}]}
    return json.dumps({"type": "Difference", "matrix": sc([1, 0, 0, 0.33, 0, 1, 0, 0.156, 0, 0, 1, 0.000000000000014, 0, 0, 0, 1], 1.0 / 60.0),
                       "children": [{"type": "cylinder", "matrix": sc([42.5, 0, 0, 0, 0, 49.6, 0, 0, 0, 0, 10, 0, 0, 0, 0, 1])}, inner]})


def simple_cone():
    """SIMPLE_CONE demo, examples/js/example_objects.js:79 (icone, matrix diag 8) scaled to 0.8."""
    return json.dumps({"type": "icone", "matrix": [0.8, 0, 0, 0, 0, 0.8, 0, 0, 0, 0, 0.8, 0, 0, 0, 0, 1]})


def bushy_tree(seed=7):
    """A balanced CSG tree (both children are CSG nodes) to exercise the SAVE/RESTORE frame stack."""
    rng = np.random.default_rng(seed)
    def prim(kind):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        return {"type": kind, "matrix": _mat(q * rng.uniform(0.3, 0.6), rng.uniform(-0.3, 0.3, size=3))}
    def aff():
        return _mat(_rot_axis(rng.normal(size=3), rng.uniform(-20, 20)) * rng.uniform(0.9, 1.1), rng.uniform(-0.05, 0.05, size=3))
    l = {"type": "Union", "matrix": aff(), "children": [prim("cube"), prim("cylinder"), prim("ellipsoid")]}
    r = {"type": "Intersection", "matrix": aff(), "children": [prim("cone"), {"type": "Difference", "matrix": aff(), "children": [prim("ellipsoid"), prim("cube")]}]}
    return json.dumps({"type": "Difference", "matrix": aff(), "children": [l, r]})


# ---- the rest of the factory's analytic vocabulary (SURVEY §8 f-N1) ----------------------------------------------------
def n1_torus():
    """itorus (object_factory.hpp:163-173): major radius 4*0.2, tube 0.2 in local units, tilted and scaled by 0.55."""
    return json.dumps({"type": "itorus", "matrix": _mat(_rot_axis([1, 0.3, 0], 25.0) * 0.55, [0.02, -0.01, 0.03])})


def n1_heart():
    """iheart (object_factory.hpp:153-162), scaled by 0.45."""
    return json.dumps({"type": "iheart", "matrix": _mat(_rot_axis([0, 0, 1], 15.0) * 0.45, [0.0, 0.0, -0.05])})


def n1_tetrahedron():
    """tetrahedron (object_factory.hpp:175-188): corners moved by the forward matrix."""
    return json.dumps({"type": "tetrahedron", "matrix": _mat(_rot_axis([1, 1, 1], 20.0) * 0.9, [0.01, 0.02, -0.03]),
                       "corners": [[-0.5, -0.4, -0.45], [0.55, -0.35, -0.4], [0.0, 0.6, -0.35], [0.05, 0.0, 0.6]]})


def n1_mushroom():
    """implicit_double_mushroom (object_factory.hpp:86-99)."""
    return json.dumps({"type": "implicit_double_mushroom", "matrix": _mat(_rot_axis([0, 1, 0], 10.0) * 1.1, [0.0, 0.03, 0.0])})


def n1_half_plane_cut():
    """Intersection[ellipsoid, half_plane] (object_factory.hpp:396-434): a sphere cut by an oblique plane."""
    hp = {"type": "half_plane", "matrix": _mat(_rot_axis([1, 0, 0], 12.0), [0.0, 0.0, 0.05]),
          "plane_vector": [0.2, 0.1, 1.0], "plane_point": [0.0, 0.0, -0.1]}
    return json.dumps({"type": "Intersection", "matrix": EYE,
                       "children": [{"type": "ellipsoid", "matrix": _mat(np.eye(3) * 1.3, [0, 0, 0])}, hp]})


def n1_lid_slab():
    """Difference[ellipsoid, top_bottom_lid] (object_factory.hpp:480-505): the lid is positive where |z| > 0.5."""
    return json.dumps({"type": "Difference", "matrix": _mat(np.eye(3) * 0.8, [0, 0, 0]),
                       "children": [{"type": "ellipsoid", "matrix": _mat(np.eye(3) * 1.6, [0, 0, 0])},
                                    {"type": "top_bottom_lid", "matrix": EYE}]})


def n1_medley():
    """Every N1 primitive in one tree: Difference[Union[torus, heart, tetrahedron, mushroom], half_plane]."""
    tor = {"type": "itorus", "matrix": _mat(_rot_axis([1, 0, 0], 90.0) * 0.5, [-0.35, 0.0, 0.0])}
    hrt = {"type": "iheart", "matrix": _mat(np.eye(3) * 0.3, [0.45, 0.0, 0.1])}
    tet = {"type": "tetrahedron", "matrix": _mat(np.eye(3) * 0.6, [0.0, 0.45, 0.0]),
           "corners": [[-0.5, -0.4, -0.45], [0.55, -0.35, -0.4], [0.0, 0.6, -0.35], [0.05, 0.0, 0.6]]}
    msh = {"type": "implicit_double_mushroom", "matrix": _mat(_rot_axis([1, 0, 0], 90.0) * 0.7, [0.0, -0.45, 0.0])}
    hp = {"type": "half_plane", "matrix": _mat(np.eye(3), [0, 0, 0]), "plane_vector": [0.0, 0.3, 1.0], "plane_point": [0.0, 0.0, 0.35]}
    u = {"type": "Union", "matrix": _mat(_rot_axis([0, 0, 1], 8.0) * 0.95, [0.01, 0.0, 0.0]), "children": [tor, hrt, tet, msh]}
    return json.dumps({"type": "Difference", "matrix": EYE, "children": [u, hp]})


def random_tree(seed, depth=4, all_prims=False):
    """A random CSG tree for property tests: random shape (chains and bushy parts mixed), every operator, every analytic
    primitive except the double-precision ones (torus / heart / mushroom agree with glibc's pow only on most points),
    n-ary unions, random similarity-ish transforms on every node."""
    rng = np.random.default_rng(seed)
    kinds = ["cube", "cylinder", "ellipsoid", "icone", "meta_balls", "tetrahedron", "half_plane", "top_bottom_lid"]
    if all_prims:      # for checks that do not compare with glibc's pow (interval bounds, masks)
        kinds += ["itorus", "iheart", "implicit_double_mushroom"]

    def aff(scale_lo, scale_hi, shift):
        R = _rot_axis(rng.normal(size=3), rng.uniform(-40, 40)) * rng.uniform(scale_lo, scale_hi)
        return _mat(R, rng.uniform(-shift, shift, size=3))

    def prim():
        k = kinds[int(rng.integers(len(kinds)))]
        node = {"type": k, "matrix": aff(0.5, 1.1, 0.3)}
        if k == "tetrahedron":
            node["corners"] = [[-0.5, -0.4, -0.45], [0.55, -0.35, -0.4], [0.0, 0.6, -0.35], [0.05, 0.0, 0.6]]
        elif k == "half_plane":
            node["plane_vector"] = [float(np.float32(v)) for v in rng.normal(size=3)]
            node["plane_point"] = [float(np.float32(v)) for v in rng.uniform(-0.2, 0.2, size=3)]
        return node

    def tree(d):
        if d == 0 or rng.uniform() < 0.25:
            return prim()
        op = ["Union", "Intersection", "Difference"][int(rng.integers(3))]
        n = int(rng.integers(2, 5)) if op == "Union" else 2      # the factory folds n-ary unions; the others take two
        return {"type": op, "matrix": aff(0.9, 1.1, 0.05), "children": [tree(d - 1) for _ in range(n)]}

    root = tree(depth)
    if "children" not in root:
        root = {"type": "Union", "matrix": EYE, "children": [root, prim()]}
    return json.dumps(root)
