"""CPU: oracle/ref_gpu_chain.py (the reference's operator sequence, timed on the B200 by bench.py's `reference_gpu`
leg) computes what the pinned oracle computes -- and, when the reference tree is present, what the reference's own
functions compute."""
import numpy as np
import pytest
import torch

from oracle import ctdet_oracle as O
from oracle import gen_golden as G
from oracle import ref_gpu_chain as RC
from oracle import ref_loader as RL


def test_chain_equals_oracle_on_cpu():
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = G.build_inputs("dense_small")
    gt_np = O.gen_heatmap_batch(bl, cl, H, W, C)
    gt = RC.target_gen_host(bl, cl, H, W, C, "cpu")
    for k in gt_np:
        assert np.array_equal(gt[k].numpy(), gt_np[k]), k
    t = [torch.from_numpy(a).clone().requires_grad_(True) for a in (hm, wh, reg)]
    torch.set_num_threads(1)
    out = RC.losses_fwd_bwd(t[0], t[1], t[2], gt, alpha=[1.0])
    ref, grads = O.losses_and_grads(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), gt_np, alpha=[1.0])
    for k in ref:
        np.testing.assert_allclose(float(out[k]), float(ref[k]), rtol=1e-6)
    for a, g in zip(t, grads):
        np.testing.assert_allclose(a.grad.numpy(), g.numpy(), rtol=1e-5, atol=1e-9)
    b, s, c = RC.decode_loop(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), 4, K)
    ob, os_, oc, _ = O.decode(O._sigmoid_clamp(hm), wh, reg, 4, K)
    assert np.array_equal(c.numpy(), oc) and np.array_equal(b.numpy(), ob)
    np.testing.assert_allclose(s.numpy(), os_, rtol=2e-7)


@pytest.mark.skipif(not RL.available(), reason="reference tree not present")
def test_chain_equals_live_reference():
    bl, cl, hm, wh, reg, (B, C, H, W, K, full) = G.build_inputs("tiny")
    gt = RC.target_gen_host(bl, cl, H, W, C, "cpu")
    t = [torch.from_numpy(a).clone().requires_grad_(True) for a in (hm, wh, reg)]
    torch.set_num_threads(1)
    out = RC.losses_fwd_bwd(t[0], t[1], t[2], gt, alpha=[1.0])
    ref, grads = RL.ref_losses(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), gt, alpha=[1.0])
    for k in ref:
        assert float(out[k]) == float(ref[k]), k                      # same ops in the same order: bit-equal
    for a, g in zip(t, grads):
        assert torch.equal(a.grad, g)
    heat = torch.clamp(torch.from_numpy(hm).clone().sigmoid_(), min=1e-4, max=1 - 1e-4)
    rb, rs, rc = RL.ref_decode(heat, torch.from_numpy(wh), torch.from_numpy(reg), 4, K)
    b, s, c = RC.decode_loop(torch.from_numpy(hm), torch.from_numpy(wh), torch.from_numpy(reg), 4, K)
    assert torch.equal(b, rb) and torch.equal(s, rs) and torch.equal(c, rc)
