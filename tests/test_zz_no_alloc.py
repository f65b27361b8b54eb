"""ss_grad_norm_noalloc (include/ssb200.h): the gradient norm for callers that cannot allocate.  It has no counterpart
in the reference; its oracle is the existing entry point ss_grad_norm,
which runs the same two kernels with a caller workspace.  The file sorts last on purpose:
a failure here must not hide any other test.

Everything goes through the C ABI (ctypes); torch only owns the input / output buffers, all created BEFORE the calls
under test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-5          # the bar tests/test_gpu_train.py holds ss_grad_norm to (relative)


def _abi():
    from ssb200._lib import LL, SSBError, call, lib, stream_ptr
    lib().ss_grad_norm_noalloc_max_n.restype = ctypes.c_longlong
    return LL, SSBError, call, lib(), stream_ptr


def test_noalloc_grad_norm_matches_workspace_entry_point_two_calls_in_a_row():
    LL, SSBError, call, lib, stream_ptr = _abi()
    assert lib.ss_grad_norm_noalloc_max_n() == 64 << 20
    torch.manual_seed(0)
    # sizes: empty, scalar tail path, exactly one block, one block + 1, unaligned, the trainer's flat buffer class (13.6 M)
    sizes = (0, 1, 7, 8192, 8193, 100003, 13_600_000 + 3)
    cases = []
    for n in sizes:
        g1 = torch.randn(n, device="cuda") * 3
        g2 = torch.randn(n, device="cuda") * 0.01 + 1.0       # a different norm: stale partials would show
        partial = torch.empty(max(1, lib.ss_optim_parts(ctypes.c_longlong(n))), device="cuda")
        refs = (float(g1.double().norm()), 0.125 * float(g2.double().norm()))      # torch allocates here, not below
        cases.append((n, g1, g2, partial, torch.full((4,), -1.0, device="cuda"), refs))
    torch.cuda.synchronize()
    st = stream_ptr()
    free_before = None
    for n, g1, g2, partial, out, (ref1, ref2) in cases:
        # two calls in a row on the shared scratch, large after small and small after large as the loop goes on
        call("grad_norm_noalloc", g1, LL(n), 1.0, out[0:], st)
        call("grad_norm_noalloc", g2, LL(n), 0.125, out[1:], st)
        call("grad_norm", g1, LL(n), 1.0, partial, out[2:], st)
        call("grad_norm", g2, LL(n), 0.125, partial, out[3:], st)
        torch.cuda.synchronize()
        if free_before is None:
            free_before = torch.cuda.mem_get_info()[0]         # after the first launches (lazy module load is behind us)
        a1, a2, b1, b2 = (float(x) for x in out.cpu())
        assert abs(a1 - b1) <= TOL * max(b1, 1e-30), (n, a1, b1)
        assert abs(a2 - b2) <= TOL * max(b2, 1e-30), (n, a2, b2)
        if n:
            assert abs(a1 - ref1) <= TOL * ref1 and abs(a2 - ref2) <= TOL * ref2, (n, a1, ref1, a2, ref2)
        else:
            assert a1 == 0.0 and a2 == 0.0
    # a smaller problem right after the largest one must not see its partials
    n, g1, _, _, out, (ref1, _) = cases[2]
    call("grad_norm_noalloc", g1, LL(n), 1.0, out[0:], st)
    torch.cuda.synchronize()
    assert abs(float(out[0]) - ref1) <= TOL * ref1
    assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was allocated by the calls"


def test_noalloc_grad_norm_rejects_sizes_beyond_the_static_scratch():
    LL, SSBError, call, lib, stream_ptr = _abi()
    out = torch.full((1,), -1.0, device="cuda")
    g = torch.ones(16, device="cuda")
    too_many = lib.ss_grad_norm_noalloc_max_n() + 1           # rejected before any memory is touched
    with pytest.raises(SSBError, match="exceeds the static scratch"):
        call("grad_norm_noalloc", g, LL(too_many), 1.0, out, stream_ptr())
    with pytest.raises(SSBError, match="bad arguments"):
        call("grad_norm_noalloc", g, LL(-1), 1.0, out, stream_ptr())
    torch.cuda.synchronize()
    assert float(out) == -1.0
    # and the library is still usable afterwards
    call("grad_norm_noalloc", g, LL(16), 1.0, out, stream_ptr())
    torch.cuda.synchronize()
    assert abs(float(out) - 4.0) <= TOL * 4.0


def test_noalloc_grad_norm_runs_under_stream_capture():
    """The use the entry point exists for: a cudaMalloc inside a capturing stream fails the capture, so a graph that
    captures two calls and replays to the workspace entry point's result shows that nothing allocates."""
    LL, SSBError, call, lib, stream_ptr = _abi()
    torch.manual_seed(1)
    n = 1_000_003
    g1, g2 = torch.randn(n, device="cuda"), torch.randn(n, device="cuda") * 2
    out = torch.zeros(4, device="cuda")
    partial = torch.empty(lib.ss_optim_parts(ctypes.c_longlong(n)), device="cuda")
    call("grad_norm_noalloc", g1, LL(n), 1.0, out[0:], stream_ptr())       # set-up (symbol address) outside the capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        call("grad_norm_noalloc", g1, LL(n), 1.0, out[0:], stream_ptr())
        call("grad_norm_noalloc", g2, LL(n), 1.0, out[1:], stream_ptr())
    out.zero_()
    graph.replay()
    call("grad_norm", g1, LL(n), 1.0, partial, out[2:], stream_ptr())
    call("grad_norm", g2, LL(n), 1.0, partial, out[3:], stream_ptr())
    torch.cuda.synchronize()
    a1, a2, b1, b2 = (float(x) for x in out.cpu())
    assert abs(a1 - b1) <= TOL * b1 and abs(a2 - b2) <= TOL * b2, (a1, b1, a2, b2)
