"""Pins the tcgen05 conventions of csrc/tc05.cuh (shared-memory descriptors, K-major / MN-major operands,
TMEM accumulator layouts for M = 128 and M = 64) against torch matmuls on one tile."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_tc_probe():
    from radnerf_b200._lib import call, stream_arg
    torch.manual_seed(0)
    dev = "cuda"
    X = torch.randn(128, 64, device=dev).half()
    W = torch.randn(64, 64, device=dev).half()
    Y = torch.randn(128, 32, device=dev).half()
    o1 = torch.zeros(128, 64, device=dev); o2 = torch.zeros(128, 32, device=dev); o3 = torch.zeros(128, 64, device=dev)
    call("rnb_tc_probe", X, W, Y, o1, o2, o3, stream_arg(torch.device(dev)))
    torch.cuda.synchronize()
    r1 = X.float() @ W.float().t()
    r3 = X.float() @ W.float()
    r2 = X.float().t() @ Y.float()         # (64, 32)
    print("err1", float((o1 - r1).abs().max()), "err3", float((o3 - r3).abs().max()))
    # discover the lane layout of the M=64 accumulator
    for name, rows in (("lanes 0-63", list(range(64))),
                       ("16 per subpartition", [32 * (i // 16) + i % 16 for i in range(64)])):
        print(name, float((o2[rows] - r2).abs().max()))
    assert float((o1 - r1).abs().max()) < 1e-2
    assert float((o3 - r3).abs().max()) < 1e-2
    rows = [32 * (i // 16) + i % 16 for i in range(64)]
    ok_a = float((o2[:64] - r2).abs().max()) < 5e-2
    ok_b = float((o2[rows] - r2).abs().max()) < 5e-2
    assert ok_a or ok_b
