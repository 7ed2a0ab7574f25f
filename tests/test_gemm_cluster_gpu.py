"""The cluster / TMA-multicast schedules of the tcgen05 W8A8 GEMM (weight tile shared by CM CTAs)
must give the same exact int32 accumulators and fp16 outputs as the oracle, including ragged M
(fewer m-tiles than the cluster is wide, so some ranks only see zero-filled rows)."""
import numpy as np
import pytest
import torch

from b200stage import ops
from oracle import w8a8 as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.mark.parametrize("sched", [10, 11, 12, 13, 14, 15, 16])
@pytest.mark.parametrize("M,N,K", [(512, 1024, 1024), (300, 512, 2048), (100, 256, 256), (1024, 4096, 512), (640, 768, 4096)])
def test_gemm_cluster_schedules_exact(sched, M, N, K):
    r = np.random.default_rng(sched * 1000 + M + N + K)
    a = r.integers(-127, 128, (M, K), dtype=np.int8)
    w = r.integers(-127, 128, (N, K), dtype=np.int8)
    sa = r.uniform(1e-3, 5e-2, M).astype(np.float32)
    sw = r.uniform(1e-4, 2e-3, N).astype(np.float32)
    bias = (0.1 * r.standard_normal(N)).astype(np.float16)
    res = r.standard_normal((M, N)).astype(np.float16)
    ref, acc_ref = O.gemm_w8a8(a, w, sa, sw, bias, res)
    acc = torch.full((M, N), -5, dtype=torch.int32, device=DEV)
    out = ops.gemm_w8a8(t(a), t(w), t(sa), t(sw), t(bias), t(res), acc_dump=acc, schedule=sched)
    torch.cuda.synchronize()
    assert np.array_equal(acc.cpu().numpy(), acc_ref)
    assert np.array_equal(out.cpu().numpy().view(np.uint16), ref.view(np.uint16))


@pytest.mark.parametrize("sched,M", [(18, 1), (18, 9), (18, 16), (18, 32), (19, 33), (19, 64), (17, 40), (0, 12), (0, 16)])
@pytest.mark.parametrize("N,K", [(1024, 1024), (4096, 4096), (22016, 512), (512, 11008)])
def test_gemm_small_batch_schedules_exact(sched, M, N, K):
    """Decode-batch schedules (only 32 / 64 activation rows staged, the MMA's other A rows alias weight bytes and
    are never stored) and the heuristic that picks them: exact accumulators and outputs, several tiles per CTA."""
    r = np.random.default_rng(sched * 100 + M + N + K)
    a = r.integers(-127, 128, (M, K), dtype=np.int8)
    w = r.integers(-127, 128, (N, K), dtype=np.int8)
    sa = r.uniform(1e-3, 5e-2, M).astype(np.float32)
    sw = r.uniform(1e-4, 2e-3, N).astype(np.float32)
    res = r.standard_normal((M, N)).astype(np.float16)
    ref, acc_ref = O.gemm_w8a8(a, w, sa, sw, None, res)
    acc = torch.full((M, N), -5, dtype=torch.int32, device=DEV)
    out = ops.gemm_w8a8(t(a), t(w), t(sa), t(sw), None, t(res), acc_dump=acc, schedule=sched)
    torch.cuda.synchronize()
    assert np.array_equal(acc.cpu().numpy(), acc_ref)
    assert np.array_equal(out.cpu().numpy().view(np.uint16), ref.view(np.uint16))


def test_small_batch_schedule_refuses_a_batch_it_cannot_stage():
    a = torch.zeros((40, 256), dtype=torch.int8, device=DEV)
    w = torch.zeros((256, 256), dtype=torch.int8, device=DEV)
    s1, s2 = torch.ones(40, device=DEV), torch.ones(256, device=DEV)
    with pytest.raises(ops._abi.B2Error):
        ops.gemm_w8a8(a, w, s1, s2, schedule=18)


def test_gemm_cluster_many_tiles_per_cluster():
    """More cluster tiles than clusters: the persistent loop and the ring phases wrap several times."""
    M, N, K = 512, 22016, 512
    g = torch.Generator(device=DEV).manual_seed(3)
    a = torch.randint(-127, 128, (M, K), generator=g, device=DEV, dtype=torch.int8)
    w = torch.randint(-127, 128, (N, K), generator=g, device=DEV, dtype=torch.int8)
    one_m, one_n = torch.ones(M, device=DEV), torch.ones(N, device=DEV)
    accs = []
    for sched in (10, 12, 13):
        acc = torch.empty((M, N), dtype=torch.int32, device=DEV)
        ops.gemm_w8a8(a, w, one_m, one_n, acc_dump=acc, schedule=sched)
        accs.append(acc)
    ref = (a[:64].double() @ w.double().T).to(torch.int32)
    assert torch.equal(accs[0][:64], ref)
    assert torch.equal(accs[0], accs[1]) and torch.equal(accs[0], accs[2])
