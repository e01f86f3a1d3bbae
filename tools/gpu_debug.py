"""GPU diagnostics: device-built BVH vs CPU-emulated build; device traversal vs CPU traversal of the same BVH."""
import ctypes as C, sys, os, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracer2_b200 as rt
from oracle import oracle_ffi as O
sys.path.insert(0, "tests")
import conftest
L = C.CDLL("tests/hostemu/libhostemu.so")
L.hostemu_trace_external.argtypes = [C.c_void_p] * 3 + [C.c_uint32, C.c_uint32, C.c_void_p]

def run(kind, scale, w, h, seed=0):
    sc = rt.Scene(kind, scale, 64, seed)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights); g.refrence_mode = 1
    ctx = rt.Renderer(0); rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h); lp.update_uniform(g)
    lp.trace_primary()
    got = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(h, w, 8)
    ref = O.OracleScene(sc.desc).trace_primary(g, w, h, flags=0)
    nodes = ctx.read(rt.BUF_BLAS_NODES, np.uint8); tris = ctx.read(rt.BUF_BLAS_TRIS, np.float32); s2p = ctx.read(rt.BUF_BLAS_SLOT_TO_PRIM, np.uint32)
    n = sc.num_triangles
    print(f"--- kind={kind} scale={scale} tris={n} nodes={len(nodes)//80} perm_ok={np.array_equal(np.sort(s2p), np.arange(n))}")
    ext = np.zeros((h, w, 8), np.uint32)
    L.hostemu_trace_external(nodes.ctypes.data_as(C.c_void_p), tris.ctypes.data_as(C.c_void_p), C.cast(C.byref(g), C.c_void_p), w, h, ext.ctypes.data_as(C.c_void_p))
    print("  device BVH traversed on CPU == oracle:", np.array_equal(ext, ref), " mismatching px:", int((ext != ref).any(axis=2).sum()))
    print("  device traversal == oracle:", np.array_equal(got, ref), " mismatching px:", int((got != ref).any(axis=2).sum()), "ids only:", int((got[:,:,:3] != ref[:,:,:3]).any(axis=2).sum()))
    bad = np.argwhere((got != ref).any(axis=2))
    for y, x in bad[:6]:
        print("   px", x, y, "gpu", got[y, x, :3], got[y, x, 3:6].view(np.float32), "ref", ref[y, x, :3], ref[y, x, 3:6].view(np.float32))
    ctx.close()

run(rt.SCENE_CORNELL, 1, 64, 64)
run(rt.SCENE_RANDOM, 1, 64, 64, 3)
run(rt.SCENE_HALL, 0.125, 160, 90)
run(rt.SCENE_HALL, 1.0, 320, 180)
