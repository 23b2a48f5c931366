"""dnrp_fec_decode_plcf_noalloc (the entry point that allocates nothing: scratch in a static __device__ array) against
dnrp_fec_decode_plcf on the same LLRs, through the C ABI, on two calls in a row that share the scratch array.
Bit exact, like every FEC test here.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from oracle import fec
from conftest import CONFIGS

pytestmark = pytest.mark.gpu
NET = 123456789


def _llrs(n, seed):
    rng = np.random.default_rng(seed)
    llr = np.zeros((n, 196), dtype=np.int16)
    for i in range(n):
        ty = 1 + (i % 2)
        plcf = rng.integers(0, 256, 5 if ty == 1 else 10, dtype=np.uint8)
        bits = fec.unpack_bits(fec.pcc_encode(plcf, ty, i % 4 == 1, i % 8 >= 4), 196).astype(np.float64)
        sigma = 10 ** (-(-4.0 + 0.25 * (i % 32)) / 20)
        y = (2.0 * bits - 1.0) + sigma * rng.standard_normal(196)
        llr[i] = np.clip(np.round(y * 2 / sigma ** 2 * 8), -127, 127).astype(np.int16)
    return llr


def test_noalloc_plcf_decode_equals_existing_entry_point_twice(product):
    import torch
    dev = torch.device("cuda:0")
    phy = product.Phy(CONFIGS["C1_loopback"], max_batch=96, network_id=NET, plcf_type=1)
    n_decoded = 0
    # second call: other inputs, smaller batch, scratch still holding the first call's soft values
    for n, seed in ((96, 21), (40, 22)):
        llr = _llrs(n, seed)
        ok_ref, plcf_ref, mask_ref, nit_ref = phy.fec_decode_plcf(llr)
        d_llr = torch.from_numpy(llr).to(dev)
        d_ok = torch.full((n, 2), 0xEE, dtype=torch.uint8, device=dev)
        d_plcf = torch.full((n, 2, 16), 0xEE, dtype=torch.uint8, device=dev)
        d_mask = torch.full((n, 2), 0x6EEEEEEE, dtype=torch.int32, device=dev)
        d_nit = torch.full((n, 2), 0xEE, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()          # the context's stream is non-blocking: inputs must have landed
        phy.fec_decode_plcf_noalloc(n, d_llr.data_ptr(), d_ok.data_ptr(), d_plcf.data_ptr(), d_mask.data_ptr(),
                                    d_nit.data_ptr())
        assert product.lib().dnrp_ctx_synchronize(phy._ctx) == 0
        ok = d_ok.cpu().numpy().astype(bool)
        assert np.array_equal(ok, ok_ref)
        assert np.array_equal(d_nit.cpu().numpy(), nit_ref)
        assert np.array_equal(d_mask.cpu().numpy().view(np.uint32), mask_ref)
        plcf = d_plcf.cpu().numpy()        # a type 1 decode writes 5 bytes, a type 2 decode 10 (rlen of the code block)
        assert np.array_equal(plcf[:, 0, :5], plcf_ref[:, 0, :5]) and np.array_equal(plcf[:, 1, :10], plcf_ref[:, 1, :10])
        n_decoded += int(ok.sum())
    assert n_decoded >= 40                # the comparison is not between two all-failed decodes
    phy.close()


def test_noalloc_rejects_batches_beyond_the_static_scratch(product):
    import torch
    phy = product.Phy(CONFIGS["C1_loopback"], max_batch=4, network_id=NET, plcf_type=1)
    n = 4096 + 1                           # DNRP_NOALLOC_MAX_PACKETS + 1
    buf = torch.zeros(n * 196, dtype=torch.int16, device="cuda:0")
    out = torch.zeros(n * 32, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(product.DnrpError, match="DNRP_NOALLOC_MAX_PACKETS"):
        phy.fec_decode_plcf_noalloc(n, buf.data_ptr(), out.data_ptr(), out.data_ptr(), out.data_ptr(), out.data_ptr())
    phy.close()
