"""rt_intersect_spheres_noalloc (csrc/noalloc.cu): the allocation-free variant of the linear-scan closest-hit kernel, whose
scratch is a static __device__ array.  No counterpart in the reference, so the
checker is the existing entry point for the same question, rt_trace_rays in strict mode: ids and hit distances must be
bit-identical (the tolerance every strict-mode intersection test in this suite uses).

Collected last on purpose (a failure here must not hide another test).  The inputs and outputs are
device memory owned by the caller -- torch tensors here; the entry point itself allocates nothing.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _spheres(seed, n, extent=30.0):
    rng = np.random.default_rng(seed)
    c = np.stack([rng.uniform(-extent, extent, n), rng.uniform(0.2, 9.0, n), rng.uniform(-extent, extent, n)], axis=1)
    r = rng.uniform(0.05, 0.9, (n, 1))
    return np.ascontiguousarray(np.concatenate([c, r], axis=1))


def _rays(seed, n, spheres, extent=35.0):
    rng = np.random.default_rng(seed)
    o = np.stack([rng.uniform(-extent, extent, n), rng.uniform(0.1, 9.0, n), rng.uniform(-extent, extent, n)], axis=1)
    d = rng.normal(size=(n, 3))
    d[: n // 2] /= np.linalg.norm(d[: n // 2], axis=1, keepdims=True)   # half unit, half un-normalised directions
    k = min(len(spheres), n // 8)
    o[:k] = spheres[:k, :3]                                             # rays that start inside a sphere (second root)
    return np.ascontiguousarray(np.concatenate([o, d], axis=1))


def _existing(gpu, spheres, rays):
    """rt_trace_rays on a scene that holds just these spheres (object i = sphere i)"""
    L = gpu.lib()
    h = L.rt_scene_new()
    try:
        d3 = lambda *v: (C.c_double * 3)(*v)   # noqa: E731
        assert L.rt_scene_add_material_phong(h, d3(.5, .5, .5), d3(0, 0, 0), C.c_double(1.0), d3(0, 0, 0)) == 0
        mats = np.zeros(len(spheres), dtype=np.uint32)
        if len(spheres):
            assert L.rt_scene_add_spheres(h, len(spheres), spheres.ctypes.data_as(C.POINTER(C.c_double)), mats.ctypes.data_as(C.POINTER(C.c_uint32))) == 0
        o = gpu.RenderOpts()
        L.rt_render_opts_default(C.byref(o))
        o.accel, o.fp_mode = gpu.ACCEL_LINEAR, gpu.FP_STRICT
        n = len(rays)
        ids, t = np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.float64)
        rc = L.rt_trace_rays(h, C.byref(o), n, rays.ctypes.data_as(C.POINTER(C.c_double)), ids.ctypes.data_as(C.POINTER(C.c_int32)),
                             t.ctypes.data_as(C.POINTER(C.c_double)), None)
        assert rc == 0, L.rt_last_error()
        return ids, t
    finally:
        L.rt_scene_free(h)


def _no_alloc(gpu, torch, d_sph, n_sph, d_rays, n_rays, d_ids, d_t):
    L = gpu.lib()
    st = torch.cuda.current_stream().cuda_stream
    rc = L.rt_intersect_spheres_noalloc(0, C.c_void_p(st), C.c_void_p(d_sph.data_ptr()), n_sph, C.c_void_p(d_rays.data_ptr()), n_rays,
                                        C.c_void_p(d_ids.data_ptr()), C.c_void_p(d_t.data_ptr()))
    assert rc == 0, L.rt_last_error()
    torch.cuda.synchronize()
    return d_ids[:n_rays].cpu().numpy(), d_t[:n_rays].cpu().numpy()


def _free_bytes(torch):
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info(0)[0]


def test_no_alloc_entry_point_matches_rt_trace_rays_on_two_calls_in_a_row(gpu):
    import torch
    torch.cuda.set_device(0)
    # call 1: several shared-memory tiles of spheres and a ragged last ray tile; call 2: FEWER spheres and other rays, so
    # anything the first call left in the static array (records past n_spheres, the cursor, the counter) would show
    cases = [(_spheres(11, 3000), 50_021), (_spheres(12, 700), 20_000)]
    cases = [(s, _rays(21 + i, n, s)) for i, (s, n) in enumerate(cases)]
    # the caller's buffers, allocated once for both calls (outputs poisoned before each call)
    d_sph = torch.empty(3000 * 4, dtype=torch.float64, device="cuda")
    d_rays = torch.empty(50_021 * 6, dtype=torch.float64, device="cuda")
    d_ids = torch.empty(50_021, dtype=torch.int32, device="cuda")
    d_t = torch.empty(50_021, dtype=torch.float64, device="cuda")
    for sph, rays in cases:
        want_ids, want_t = _existing(gpu, sph, rays)
        d_sph[: sph.size].copy_(torch.from_numpy(sph.reshape(-1)))
        d_rays[: rays.size].copy_(torch.from_numpy(rays.reshape(-1)))
        d_ids.fill_(-7)
        d_t.fill_(float("nan"))
        ids, t = _no_alloc(gpu, torch, d_sph, len(sph), d_rays, len(rays), d_ids, d_t)
        assert np.array_equal(ids, want_ids)
        assert np.array_equal(t, want_t), "hit distances must be bit-identical to rt_trace_rays (strict)"
        assert (ids >= 0).mean() > 0.05 and (ids < 0).any() and np.isinf(t[ids < 0]).all()
    # the same call again, unchanged inputs: same answer (the array keeps the last call's contents)
    # -- and the device's free memory does not move across it (the module and its static array were loaded by the calls above)
    free_before = _free_bytes(torch)
    ids2, t2 = _no_alloc(gpu, torch, d_sph, len(sph), d_rays, len(rays), d_ids, d_t)
    assert _free_bytes(torch) >= free_before, "the no-alloc entry point took device memory"
    assert np.array_equal(ids2, want_ids) and np.array_equal(t2, want_t)


def test_no_alloc_entry_point_edges_and_limit(gpu):
    import torch
    torch.cuda.set_device(0)
    L = gpu.lib()
    sph = _spheres(5, 64)
    rays = _rays(6, 1000, sph)
    d_sph = torch.from_numpy(sph.reshape(-1)).cuda()
    d_rays = torch.from_numpy(rays.reshape(-1)).cuda()
    d_ids = torch.full((1000,), -7, dtype=torch.int32, device="cuda")
    d_t = torch.zeros(1000, dtype=torch.float64, device="cuda")
    # beyond the static scratch: a clear error, nothing launched, outputs untouched
    rc = L.rt_intersect_spheres_noalloc(0, None, C.c_void_p(d_sph.data_ptr()), gpu.NOALLOC_MAX_SPHERES + 1, C.c_void_p(d_rays.data_ptr()), 1000,
                                        C.c_void_p(d_ids.data_ptr()), C.c_void_p(d_t.data_ptr()))
    assert rc == -1
    msg = L.rt_last_error().decode()
    assert "RT_NOALLOC_MAX_SPHERES" in msg and str(gpu.NOALLOC_MAX_SPHERES) in msg
    torch.cuda.synchronize()
    assert (d_ids == -7).all()
    # no spheres: every ray misses
    ids, t = _no_alloc(gpu, torch, d_sph, 0, d_rays, 1000, d_ids, d_t)
    assert (ids == -1).all() and np.isinf(t).all()
    # no rays: nothing to do
    assert L.rt_intersect_spheres_noalloc(0, None, C.c_void_p(d_sph.data_ptr()), 64, C.c_void_p(d_rays.data_ptr()), 0, None, None) == 0
    # and a normal call after the rejected one still agrees with the existing entry point
    want_ids, want_t = _existing(gpu, sph, rays)
    ids, t = _no_alloc(gpu, torch, d_sph, 64, d_rays, 1000, d_ids, d_t)
    assert np.array_equal(ids, want_ids) and np.array_equal(t, want_t)
