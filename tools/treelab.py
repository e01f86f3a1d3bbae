"""Tree-quality lab (measurement only; VERDICT r1 task 6).  Builds the product's CWBVH on the CPU (tools/treelab/treelab.cu: the
product's own builder + traversal functions) with 0..k passes of insertion-based SAH optimisation (Bittner et al. 2013) on the PLOC
BVH2, and reports node steps and triangle tests per ray for (a) primary rays and (b) incoherent bounce rays (origins on the
surfaces, area-weighted; cosine-distributed directions about the facing normal) — the rays K1's bounces trace.
usage: python tools/treelab.py [--scene hall|manylight|cornell] [--scale S] [--rays N] [--passes 0,1,2,4]"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import argparse, ctypes as C, subprocess, time
import numpy as np
import raytracer2_b200 as rt

HERE = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "treelab")


def lib():
    so, src = _os.path.join(HERE, "libtreelab.so"), _os.path.join(HERE, "treelab.cu")
    deps = [src] + [_os.path.join(_os.path.dirname(HERE), "..", "raytracer2_b200", "csrc", f) for f in ("rt2_build.cuh", "rt2_bvh.cuh", "rt2_math.cuh")]
    if not _os.path.exists(so) or any(_os.path.getmtime(p) > _os.path.getmtime(so) for p in deps):
        subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-O2", "-std=c++17", "-fmad=false", "-Wno-deprecated-gpu-targets", "-diag-suppress", "20011,20014",
                               "-Xcompiler", "-fPIC,-ffp-contract=off,-mfma,-Wno-unknown-pragmas", "-shared", "-o", so, src])
    L = C.CDLL(so)
    L.treelab_build.restype = C.c_void_p
    L.treelab_build.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.treelab_build_soup.restype = C.c_void_p
    L.treelab_build_soup.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_float, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.treelab_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_float, C.c_float, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.treelab_destroy.argtypes = [C.c_void_p]
    return L


def world_triangles(sc):
    d = sc.desc.contents
    nv = d.num_vertices
    stride = C.sizeof(rt._ffi.Vertex)
    raw = np.ctypeslib.as_array(C.cast(d.vertices, C.POINTER(C.c_uint8)), shape=(nv * stride,)).view(np.float32).reshape(nv, stride // 4)
    pos = raw[:, 0:3]
    idx = np.ctypeslib.as_array(C.cast(d.indices, C.POINTER(C.c_uint32)), shape=(d.num_indices,))
    counts = np.ctypeslib.as_array(C.cast(d.index_counts, C.POINTER(C.c_uint32)), shape=(d.num_geometries,))
    tris = []
    for g in range(d.num_geometries):
        gi = d.geometries[g]
        n = int(counts[g])
        ii = idx[gi.index_offset:gi.index_offset + n].astype(np.int64) + gi.vertex_offset
        M = np.array(list(gi.transform), np.float64).reshape(4, 4).T           # column-major -> rows
        p = pos[ii].astype(np.float64)
        pw = p @ M[:3, :3].T + M[:3, 3]
        tris.append(pw.reshape(-1, 3, 3))
    tris = np.concatenate(tris)
    if d.num_instances:                                                          # flattened instancing: every (instance, triangle) pair in world space
        m = np.ctypeslib.as_array(C.cast(d.instance_transforms, C.POINTER(C.c_float)), shape=(d.num_instances, 3, 4)).astype(np.float64)
        tris = np.concatenate([tris @ m[i, :, :3].T + m[i, :, 3] for i in range(d.num_instances)])
    return tris


def bounce_rays(tris, n, seed=1):
    rng = np.random.default_rng(seed)
    e1, e2 = tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0]
    nrm = np.cross(e1, e2); area = 0.5 * np.linalg.norm(nrm, axis=1)
    pick = rng.choice(len(tris), n, p=area / area.sum())
    r1, r2 = rng.random(n), rng.random(n)
    s = np.sqrt(r1); b0, b1, b2 = 1 - s, s * (1 - r2), s * r2
    o = tris[pick, 0] * b0[:, None] + tris[pick, 1] * b1[:, None] + tris[pick, 2] * b2[:, None]
    nn = nrm[pick] / np.maximum(np.linalg.norm(nrm[pick], axis=1, keepdims=True), 1e-30)
    nn *= np.where(rng.random(n) < 0.5, 1.0, -1.0)[:, None]                       # either side (closed rooms: one of them is inside)
    u1, u2 = rng.random(n), rng.random(n)
    r, phi = np.sqrt(u1), 2 * np.pi * u2
    a = np.where(np.abs(nn[:, 0:1]) > 0.9, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
    t = np.cross(nn, a); t /= np.linalg.norm(t, axis=1, keepdims=True); b = np.cross(nn, t)
    d = t * (r * np.cos(phi))[:, None] + b * (r * np.sin(phi))[:, None] + nn * np.sqrt(np.maximum(0, 1 - u1))[:, None]
    o = o + nn * 1e-3
    return np.ascontiguousarray(np.concatenate([o, d], axis=1), np.float32)


def primary_rays(sc, w, h):
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    out = np.zeros((h, w, 6), np.float32)
    from oracle import oracle_ffi as O
    L = O.lib()
    o = np.zeros(3, np.float32); d = np.zeros(3, np.float32)
    for y in range(h):
        for x in range(w):
            L.orc_primary_ray(C.cast(C.byref(g), C.c_void_p), x, y, o.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p))
            out[y, x, :3] = o; out[y, x, 3:] = d
    return out.reshape(-1, 6)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="hall"); ap.add_argument("--scale", type=float, default=0.25)
    ap.add_argument("--rays", type=int, default=200000); ap.add_argument("--passes", default="0,1,2,4")
    ap.add_argument("--fraction", type=float, default=1.0)
    ap.add_argument("--collapse", default="0", help="comma list of collapse modes (0 = product: largest area first)")
    a = ap.parse_args()
    kind = {"hall": rt.SCENE_HALL, "manylight": rt.SCENE_MANYLIGHT, "cornell": rt.SCENE_CORNELL, "random": rt.SCENE_RANDOM, "instanced": rt.SCENE_INSTANCED}[a.scene]
    sc = rt.Scene(kind, a.scale, 16, 0)
    tris = world_triangles(sc)
    print(f"scene {a.scene} scale {a.scale}: {len(tris)} triangles")
    br = bounce_rays(tris, a.rays)
    pr = primary_rays(sc, 160, 90)
    L = lib()
    base = None
    for passes in [int(p) for p in a.passes.split(",")]:
      for cmode in [int(c) for c in a.collapse.split(",")]:
        sb, sa, wn, dp = C.c_double(), C.c_double(), C.c_uint32(), C.c_uint32()
        t0 = time.time()
        if sc.desc.contents.num_instances:
            soup = np.ascontiguousarray(tris.reshape(-1, 9), np.float32)
            hnd = L.treelab_build_soup(soup.ctypes.data_as(C.c_void_p), len(soup), passes, a.fraction, cmode, C.byref(sb), C.byref(sa), C.byref(wn), C.byref(dp))
        else:
            hnd = L.treelab_build(C.cast(sc.desc, C.c_void_p), passes, a.fraction, cmode, C.byref(sb), C.byref(sa), C.byref(wn), C.byref(dp))
        tb = time.time() - t0
        assert hnd
        row = [f"passes {passes} collapse {cmode}: SAH(node) {sb.value:.2f} -> {sa.value:.2f}, {wn.value} wide nodes, depth {dp.value}, build {tb:.1f} s"]
        res = {}
        for name, rays in (("primary", pr), ("bounce", br)):
            ns, tt = C.c_uint64(), C.c_uint64()
            t = np.zeros(len(rays), np.float32)
            L.treelab_trace(hnd, rays.ctypes.data_as(C.c_void_p), len(rays), 1e-3, 1e5, t.ctypes.data_as(C.c_void_p), C.byref(ns), C.byref(tt))
            res[name] = (ns.value / len(rays), tt.value / len(rays), float(np.isfinite(t).mean()), t)
            row.append(f"{name}: {res[name][0]:.2f} node steps, {res[name][1]:.2f} tri tests per ray (hit {res[name][2]:.3f})")
        if base is None:
            base = res
        else:
            for name in res:
                same = np.array_equal(res[name][3], base[name][3])
                row.append(f"{name} vs first: node steps {res[name][0] / base[name][0] - 1:+.1%}, tri tests {res[name][1] / base[name][1] - 1:+.1%}, same hits {same}")
        print("\n  ".join(row), flush=True)
        L.treelab_destroy(hnd)


if __name__ == "__main__":
    main()
