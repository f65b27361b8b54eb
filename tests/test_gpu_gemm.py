"""tcgen05 split-bf16 GEMM, split / transpose / column-sum kernels vs fp64 torch on the same inputs."""
import pytest
import torch

from ssb200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (200, 136, 257), (1000, 514, 1200), (2400, 600, 1999), (64, 4800, 264)])
def test_gemm_matches_fp64(M, N, K):
    a, b = _rand((M, K), 1), _rand((N, K), 2, 0.05)
    bias = _rand((N,), 3)
    A, _ = ops.split(a)
    B, _ = ops.split(b)
    out = ops.gemm(A, B, bias=bias)
    ref = a.double() @ b.double().t() + bias.double()
    err = (out.double() - ref).norm() / ref.norm()
    assert float(err) <= 3e-5, float(err)
    assert float((out.double() - ref).abs().max()) <= 1e-4 * float(ref.abs().max())
    # accumulate form: out = 0.5*A@B^T + 2*out
    out2 = out.clone()
    ops.gemm(A, B, out=out2, alpha=0.5, beta=2.0)
    ref2 = 0.5 * (a.double() @ b.double().t()) + 2.0 * out.double()
    assert float((out2.double() - ref2).norm() / ref2.norm()) <= 3e-5


def test_gemm_transposed_operands_and_rowmap():
    H, I, R = 24, 40, 300        # dW[4H, I] = dG^T @ X with dG in gate-interleaved order
    dg, x = _rand((R, 4 * H), 4), _rand((R, I), 5)
    _, dgT = ops.split(dg, want=False, want_t=True)
    _, xT = ops.split(x, want=False, want_t=True)
    out = ops.gemm(dgT, xT, rowmap=1, hidden=H)
    ref_perm = dg.double().t() @ x.double()                   # row 4u+g
    ref = ref_perm.view(H, 4, I).permute(1, 0, 2).reshape(4 * H, I)   # row g*H+u
    assert float((out.double() - ref).norm() / ref.norm()) <= 3e-5


@pytest.mark.parametrize("R,M,N", [(64, 128, 128), (300, 96, 72), (1000, 2400, 257), (4100, 514, 1200)])
def test_gemm_tn_matches_fp64(R, M, N):
    """dW = A^T B with the reduction index as the ROW of both operands (MN-major tensor-core operands, no transposes)."""
    a, b = _rand((R, M), 11), _rand((R, N), 12, 0.3)
    A, _ = ops.split(a)
    B, _ = ops.split(b)
    out, = ops.gemm_tn(A, B, M, N)
    ref = a.double().t() @ b.double()
    assert float((out.double() - ref).norm() / ref.norm()) <= 3e-5
    assert float((out.double() - ref).abs().max()) <= 1e-4 * float(ref.abs().max())


def test_gemm_tn_batched_shifted_chunked_rowmap():
    """The dW_hh form: two directions in one launch (column halves of both operands), the carried-state operand read
    with a row shift of -/+ Bt, the sum accumulated over row-block chunks (direction 0 from the top block down), gate
    rows un-interleaved on the way out."""
    Hh, Bt, T = 24, 8, 50
    R = T * Bt
    dg, o = _rand((R, 8 * Hh), 13), _rand((R, 2 * Hh), 14)
    DG, _ = ops.split(dg)
    O, _ = ops.split(o)
    outs = [torch.zeros(4 * Hh, Hh, device=DEV) for _ in range(2)]
    nblk = (R + 63) // 64
    sizes, done = [3, 2, nblk - 5], 0
    for i, n in enumerate(sizes):
        ops.gemm_tn(DG, O, 4 * Hh, Hh, outs=outs, a_col_batch=4 * Hh, b_col_batch=Hh, blk0=(nblk - done - n, done), n_blk=n,
                    shifts=(-Bt, Bt), batch=2, beta=0.0 if i == 0 else 1.0, rowmap=1, hidden=Hh)
        done += n
    od = o.double()
    for z, sh in enumerate((-Bt, Bt)):
        prev = torch.zeros(R, Hh, dtype=torch.float64, device=DEV)
        if sh < 0:
            prev[Bt:] = od[:R - Bt, z * Hh:(z + 1) * Hh]
        else:
            prev[:R - Bt] = od[Bt:, z * Hh:(z + 1) * Hh]
        ref = dg.double()[:, z * 4 * Hh:(z + 1) * 4 * Hh].t() @ prev            # rows 4u+g
        ref = ref.view(Hh, 4, Hh).permute(1, 0, 2).reshape(4 * Hh, Hh)          # rows g*H+u
        assert float((outs[z].double() - ref).norm() / ref.norm()) <= 3e-5, z


def test_split_rowperm_and_colsum():
    H, C = 600, 257
    w = _rand((4 * H, C), 6)
    scale = _rand((C,), 7)
    W, WT = ops.split(w, rowperm=1, hidden=H, col_scale=scale, want=True, want_t=True)
    ref = (w * scale).view(4, H, C).permute(1, 0, 2).reshape(4 * H, C)
    assert float((W.float() - ref).abs().max()) <= 2e-5 * float(ref.abs().max())
    assert float((WT.float() - ref.t()).abs().max()) <= 2e-5 * float(ref.abs().max())
    cs = ops.colsum(W)
    assert float((cs.double() - ref.double().sum(0)).abs().max()) <= 1e-3
    cs2 = ops.colsum(WT, rowmap=1, hidden=H)                   # columns are interleaved gate rows -> natural
    assert float((cs2.double() - (w * scale).double().sum(1)).abs().max()) <= 1e-3
