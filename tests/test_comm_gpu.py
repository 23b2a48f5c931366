"""The multi-GPU counter reduction behind the C ABI (dnrp_comm_*, dnrp_counters_allreduce): NCCL bound at
run time.  One GPU: a communicator of one rank (the reduction is the identity, but the whole path -- dlopen, unique id,
ncclCommInitRank, staging copies, ncclAllReduce on the context's stream -- runs).  Two or more GPUs: one context per GPU
inside this process, one NCCL group per reduction."""
import numpy as np
import pytest
from conftest import load_product

pytestmark = pytest.mark.gpu
CFG = (1, 1, 1, 1, 0, 2, 2048)


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def test_single_rank_communicator_is_identity_and_limits():
    product = load_product()
    phy = product.Phy(CFG, max_batch=4)
    with pytest.raises(product.DnrpError):      # no communicator yet
        phy.counters_allreduce(np.arange(4))
    uid = product.comm_unique_id()
    assert len(uid) == 128 and any(uid)
    phy.comm_init_rank(1, 0, uid)
    v = np.array([1, -5, 2 ** 40, 0, 7], dtype=np.int64)
    for op in ("sum", "min", "max"):
        assert np.array_equal(phy.counters_allreduce(v, op), v)
    with pytest.raises(product.DnrpError):      # a second communicator on the same context
        phy.comm_init_rank(1, 0, product.comm_unique_id())
    with pytest.raises(product.DnrpError):      # more counters than the staging buffer holds
        phy.counters_allreduce(np.zeros(65537, dtype=np.int64))
    big = np.arange(65536, dtype=np.int64)
    assert np.array_equal(phy.counters_allreduce(big), big)
    phy.close()                                 # destroys the communicator with the context


def test_all_gpus_of_this_process():
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs two GPUs in one process")
    product = load_product()
    phys = [product.Phy(CFG, max_batch=4, device=i) for i in range(n)]
    product.comm_init_all(phys)
    rng = np.random.default_rng(5)
    c = rng.integers(-10 ** 12, 10 ** 12, (n, 11), dtype=np.int64)
    for op, ref in (("sum", c.sum(0)), ("min", c.min(0)), ("max", c.max(0))):
        out = product.counters_allreduce_all(phys, c, op)
        for i in range(n):
            assert np.array_equal(out[i], ref), (op, i)
    for p in phys:
        p.close()
