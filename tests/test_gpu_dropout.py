"""GPU parity of training-mode dropout (reference networks/deep_sdf_decoder.py:103-104) fused into
the forward epilogues with a counter-based Philox4x32-10 stream: the keep bits are bit-exact with
the numpy oracle (bookkeeping), and the fused step's outputs / gradients match the fp32 oracle fed
with exactly those keep masks (SDF <= 2e-3, gradient cosine >= 0.999)."""
import ctypes as C

import numpy as np
import pytest
import torch

import dsdf_b200
from common import cos
from dsdf_b200 import _lib, synth
from oracle import deepsdf, philox
from oracle import train as otrain

pytestmark = pytest.mark.gpu
DROP = [0, 1, 2, 3, 4, 5, 6, 7]
KW = dict(dims=[512] * 8, dropout_prob=0.2, latent_in=[4], xyz_in_all=False, use_tanh=False,
          latent_dropout=False, weight_norm=True, norm_layers=list(range(8)))


def keep_from_lib(dec, layer, row_lo, n):
    out = torch.empty(n, 512, dtype=torch.uint8, device="cuda")
    h = dec.handle()
    _lib.check(_lib.lib().dsdf_dropout_keep(h, layer, row_lo, n, out.data_ptr(), None), h)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_keep_bits_bit_exact_with_the_philox_oracle():
    dec = dsdf_b200.Decoder(256, dropout=DROP, **KW).cuda()
    h = dec.handle()
    _lib.check(_lib.lib().dsdf_set_dropout(h, 0xFF, 0.2, (5 << 32) | 9, 77, 1 << 33), h)
    for layer, lo in [(0, 0), (3, 12345), (7, 999)]:
        got = keep_from_lib(dec, layer, lo, 300)
        want = philox.dropout_keep((5 << 32) | 9, 77, layer, (1 << 33) + lo, 300, 0.2)
        assert np.array_equal(got, want)
    _lib.check(_lib.lib().dsdf_set_dropout(h, 0x0F, 0.2, 1, 1, 0), h)
    assert keep_from_lib(dec, 6, 0, 16).all()            # layer outside the dropout list: all kept


@pytest.mark.parametrize("drop", [DROP, [1, 7]])
def test_train_step_with_dropout_matches_oracle_under_the_same_masks(drop):
    torch.manual_seed(0)
    dec = dsdf_b200.Decoder(256, dropout=drop, **KW)
    params = {k: v.detach().clone() for k, v in dec.state_dict().items()}
    for k in params:
        if k.endswith("weight_g"):
            params[k] *= 1.5
    dec.load_state_dict(params)
    dec = dec.cuda().train()
    spec = deepsdf.DecoderSpec(norm_layers=tuple(range(8)), weight_norm=True, dropout=tuple(drop), dropout_prob=0.2)
    S, B, P = 6, 3, 256
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=None)
    orc = otrain.OracleTrainer(spec, cfg, S, seed=0, device="cuda", params=params)
    tr = dsdf_b200.FusedTrainer(dec, S, dsdf_b200.StepConfig(code_bound=1.0, grad_clip=None, dropout_seed=1234),
                                latents=orc.lat_vecs.weight.data.clone())
    data = synth.scene_samples(S, P, seed=2).cuda()
    ids = torch.tensor([4, 0, 2])
    pred = tr.forward_backward(data[ids], ids.cuda(), 50)
    torch.cuda.synchronize()
    keep = {}
    for l in drop:
        k = keep_from_lib(dec, l, 0, B * P)
        keep[l] = torch.from_numpy(k[:, :253] if l == 3 else k).cuda()
        assert abs(float(keep[l].float().mean()) - 0.8) < 0.01
    orc.opt.zero_grad()
    sl, rl, pred_ref = orc.forward_loss(data[ids], ids, 50, dropout_keep=keep)
    err = (pred.clamp(-0.1, 0.1)[:, None] - pred_ref).abs().max().item()
    assert err <= 2e-3, err
    ref_flat = torch.cat([orc.params[n].grad.flatten() for n, _ in dec.named_parameters()])
    assert cos(tr.grads, ref_flat) >= 0.999
    assert cos(tr.lat_grad, orc.latent_grad()) >= 0.999
    assert float(tr.loss_buf[0]) == pytest.approx(sl, rel=1e-2)
    # eval mode: no dropout, and a different step offset gives a different mask
    p1 = pred.clone()
    tr.step_count += 1
    p2 = tr.forward_backward(data[ids], ids.cuda(), 50).clone()
    assert not torch.equal(p1, p2)
    tr.step_count -= 1
    p3 = tr.forward_backward(data[ids], ids.cuda(), 50)
    assert torch.equal(p1, p3)                            # same (seed, offset) -> bit-identical
    dec.eval()
    p4 = tr.forward_backward(data[ids], ids.cuda(), 50)
    sl0, _, pred0 = orc.forward_loss(data[ids], ids, 50, dropout_keep={l: torch.ones_like(keep[l]) for l in drop})
    assert not torch.equal(p4, p1)


def test_decoder_forward_in_train_mode_with_dropout_points_to_the_fused_step():
    dec = dsdf_b200.Decoder(256, dropout=DROP, **KW).cuda().train()
    with pytest.raises(NotImplementedError):
        dec(torch.zeros(4, 259, device="cuda"))
    dec.eval()
    assert dec(torch.zeros(4, 259, device="cuda")).shape == (4, 1)
