"""psb_push_apply_noalloc vs psb_push_apply, straight through the C ABI.

The no-alloc entry point is for callers that may not allocate device memory themselves (an
optimizer hook inside someone else's memory pool, a plug-in host that forbids allocation
after start-up): the kernel's control words and its tables come out of a static array in
the library image instead of buffers the caller would have to cudaMalloc.  The result must
be the same as the ordinary entry point's.

Two calls are made in a row on purpose: the static array is shared by every call and keeps
the previous call's contents, so the second call only gives the right answer if the entry
point clears what has to start at zero and nothing else.

This file is collected last (``pytest --collect-only``): a failure here cannot hide a
failure anywhere else in a ``-x`` run.  It passes alone in a fresh process and after the
rest of the suite (it owns every buffer it touches and uses only world == 1).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _state(seed):
    """Three variables: SGD vector (scalar path), RMSProp matrix, Adam matrix (vector path)."""
    g = torch.Generator().manual_seed(seed)
    shapes = [(10,), (24, 12), (16, 40)]
    st = {"grads": [], "master": [], "s0": [], "s1": []}
    for shp in shapes:
        n = 1
        for d in shp:
            n *= d
        st["master"].append(torch.randn(n, generator=g).cuda())
        st["s0"].append((torch.rand(n, generator=g) + 0.5).cuda())
        st["s1"].append(torch.rand(n, generator=g).cuda() * 0.1)
    return shapes, st


def _build(shapes, st, grads_flat, arena, farena):
    from psb200.store.nvl import PushSeg
    kinds = [0, 1, 2]
    hyper = [(0.1, 0, 0, 0), (0.01, 0.9, 0.0, 1e-10), (2e-4, 0.9, 0.999, 1e-8)]
    segs = (PushSeg * 3)()
    goff, aoff, foff = 0, 0, 0
    beta = torch.tensor([0.9, 0.999], device="cuda")
    for i, shp in enumerate(shapes):
        n = st["master"][i].numel()
        s = segs[i]
        s.grad_offset, s.numel = goff, n
        s.cols = shp[1] if len(shp) == 2 else n
        s.opt_kind = kinds[i]
        s.lr, s.h1, s.h2, s.eps = hyper[i]
        s.master = st["master"][i].data_ptr()
        s.slot0 = st["s0"][i].data_ptr() if kinds[i] else None
        s.slot1 = st["s1"][i].data_ptr() if kinds[i] else None
        s.beta_pow = beta.data_ptr() if kinds[i] == 2 else None
        if len(shp) == 2:
            rows, cols = shp
            ld_w, ld_wt = (cols + 7) // 8 * 8, (rows + 7) // 8 * 8
            s.w_off, s.ld_w = aoff, ld_w
            aoff += (rows * ld_w + 127) // 128 * 128
            s.wt_off, s.ld_wt = aoff, ld_wt
            aoff += (cols * ld_wt + 127) // 128 * 128
            s.f_off = -1
        else:
            s.w_off = s.wt_off = -1
            s.f_off = foff
            foff += (n + 3) // 4 * 4
        s.home, s.bcast = 0, 1
        goff += (n + 3) // 4 * 4
    assert goff <= grads_flat.numel() and aoff <= arena.numel() and foff <= farena.numel()
    return segs, beta


def _desc(beta, gstep):
    from psb200.store.nvl import PushDesc
    d = PushDesc()
    d.world, d.rank, d.n_arenas = 1, 0, 1
    d.inv_replicas = 1.0
    d.step_id = 0
    d.nbeta = 1
    d.beta[0].beta_pow = beta.data_ptr()
    d.beta[0].beta1, d.beta[0].beta2 = 0.9, 0.999
    d.global_step = gstep.data_ptr()
    d.step_delta = 1
    d.blocks_per_seg = 4
    return d


def test_noalloc_entry_matches_regular_entry_twice():
    from psb200.ops import _lib
    L = _lib.lib()
    import psb200.store.nvl as nvl  # registers psb_push_apply's signature
    fn = L.psb_push_apply_noalloc
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                   ctypes.c_void_p, ctypes.POINTER(nvl.PushDesc), ctypes.c_void_p]
    stream = torch.cuda.current_stream().cuda_stream

    def side():
        shapes, st = _state(11)
        grads = torch.zeros(4096, device="cuda")
        arena = torch.zeros(8192, dtype=torch.bfloat16, device="cuda")
        farena = torch.zeros(64, device="cuda")
        gstep = torch.zeros(1, dtype=torch.int32, device="cuda")
        segs, beta = _build(shapes, st, grads, arena, farena)
        return dict(shapes=shapes, st=st, grads=grads, arena=arena, farena=farena, gstep=gstep,
                    segs=segs, beta=beta)

    A, B = side(), side()
    # regular entry point: caller-allocated device tables and control words
    segs_dev = torch.frombuffer(bytearray(bytes(A["segs"])), dtype=torch.uint8).cuda()
    t_grads = torch.tensor([A["grads"].data_ptr()], dtype=torch.int64, device="cuda")
    t_arena = torch.tensor([A["arena"].data_ptr()], dtype=torch.int64, device="cuda")
    t_farena = torch.tensor([A["farena"].data_ptr()], dtype=torch.int64, device="cuda")
    ctrl = torch.zeros(64, dtype=torch.int32, device="cuda")    # PSB_PUSH_CTRL_WORDS
    # no-alloc entry point: host tables only
    P = ctypes.c_void_p * 1
    h_grads, h_arena, h_farena = P(B["grads"].data_ptr()), P(B["arena"].data_ptr()), P(B["farena"].data_ptr())

    for call in range(2):
        g = torch.randn(4096, generator=torch.Generator().manual_seed(100 + call)).cuda()
        A["grads"].copy_(g)
        B["grads"].copy_(g)
        dA = _desc(A["beta"], A["gstep"])
        dA.segs, dA.nseg = segs_dev.data_ptr(), 3
        dA.grad_ptrs, dA.arena_ptrs, dA.farena_ptrs = t_grads.data_ptr(), t_arena.data_ptr(), t_farena.data_ptr()
        dA.local_ctrl = ctrl.data_ptr()
        _lib.check(L.psb_push_apply(ctypes.byref(dA), stream), "psb_push_apply")
        dB = _desc(B["beta"], B["gstep"])
        _lib.check(fn(B["segs"], 3, h_grads, h_arena, h_farena, None, ctypes.byref(dB), stream),
                   "psb_push_apply_noalloc")
        torch.cuda.synchronize()
        for key in ("master", "s0", "s1"):
            for a, b in zip(A["st"][key], B["st"][key]):
                assert torch.allclose(a, b, rtol=2e-4, atol=1e-7), (call, key)
        assert torch.equal(A["arena"], B["arena"]) and torch.equal(A["farena"], B["farena"]), call
        assert torch.allclose(A["beta"], B["beta"]) and int(A["gstep"]) == int(B["gstep"]) == call + 1
    # the update did something: Adam's beta powers advanced twice
    assert abs(float(B["beta"][0]) - 0.9 ** 3) < 1e-6


def test_noalloc_entry_takes_any_segment_count_up_to_its_limit():
    """40 segments with the default number of blocks per segment: the static array holds up to 64
    segment records, and no other limit may turn a request of that size away (an earlier version
    refused it with a grid-size check that only multi-GPU launches need)."""
    from psb200.ops import _lib
    import psb200.store.nvl as nvl
    L = _lib.lib()
    fn = L.psb_push_apply_noalloc
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                   ctypes.c_void_p, ctypes.POINTER(nvl.PushDesc), ctypes.c_void_p]
    nseg, n = 40, 52
    g = torch.Generator().manual_seed(3)
    master = torch.randn(nseg * n, generator=g).cuda()
    want = master.clone()
    grads = torch.randn(nseg * n, generator=g).cuda()
    farena = torch.zeros(nseg * n, device="cuda")
    arena = torch.zeros(128, dtype=torch.bfloat16, device="cuda")
    gstep = torch.zeros(1, dtype=torch.int32, device="cuda")
    segs = (nvl.PushSeg * nseg)()
    for i in range(nseg):
        s = segs[i]
        s.grad_offset, s.numel, s.cols = i * n, n, n
        s.opt_kind, s.lr = 0, 0.5
        s.master = master.data_ptr() + 4 * i * n
        s.w_off = s.wt_off = -1
        s.f_off = i * n
        s.home, s.bcast = 0, 1
    d = nvl.PushDesc()
    d.world, d.rank, d.n_arenas, d.inv_replicas = 1, 0, 1, 1.0
    d.global_step, d.step_delta = gstep.data_ptr(), 1
    d.blocks_per_seg = 0                                  # default
    P = ctypes.c_void_p * 1
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):
        _lib.check(fn(segs, nseg, P(grads.data_ptr()), P(arena.data_ptr()), P(farena.data_ptr()), None,
                      ctypes.byref(d), stream), "psb_push_apply_noalloc")
        want -= 0.5 * grads
        torch.cuda.synchronize()
        assert torch.allclose(master, want, rtol=2e-4, atol=1e-7), call
        assert torch.equal(farena, master) and int(gstep) == call + 1


def test_noalloc_entry_rejects_oversized_requests():
    from psb200.ops import _lib
    import psb200.store.nvl as nvl
    L = _lib.lib()
    fn = L.psb_push_apply_noalloc
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                   ctypes.c_void_p, ctypes.POINTER(nvl.PushDesc), ctypes.c_void_p]
    d = nvl.PushDesc()
    d.world, d.rank, d.inv_replicas = 1, 0, 1.0
    segs = (nvl.PushSeg * 65)()
    P = ctypes.c_void_p * 1
    rc = fn(segs, 65, P(0), P(0), P(0), None, ctypes.byref(d), None)
    assert rc != 0 and b"exceed the static scratch" in L.psb_last_error()
    d.world = 17
    rc = fn(segs, 1, P(0), P(0), P(0), None, ctypes.byref(d), None)
    assert rc != 0 and b"exceed the static scratch" in L.psb_last_error()
