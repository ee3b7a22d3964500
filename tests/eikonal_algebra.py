"""The algebra of the fused Eikonal step, in plain fp32 torch WITHOUT double backward (test infrastructure).

The CUDA path (csrc: k_forward3 -> k_backward2<EIK> -> k_forward3<TAN> -> k_wgrad2 + k_eik_corr) computes the
gradient of

    L = sum|clamp(t) - gt| / N  +  lam/N * sum_rows (1 - ||G||)^2 ,   t = tanh(s),  G = d clamp(t) / d xyz

(train_deep_sdf.py:604-606,724-729) for a ReLU MLP without autograd-of-autograd.  With the ReLU masks M_l of the
forward pass held fixed (their derivative is zero almost everywhere, which is also what torch's double backward
uses), per row:

    sigma = 1[|t| <= delta] (1 - t^2)             G = sigma J,  J = ds/dxyz = W0x^T u_0 + W4x^T u_4
    g1    = sigma sign(clamp t - gt) / N          dY_l = g1 u_l   (the ordinary dgrad chain, u_l = ds/da_l)
    coef  = 2 lam / N (1 - 1/||G||)               q = dE/dG = coef G
    eps   = -2 t coef ||G||^2 / g1                (the d(1 - t^2)/ds path: extra upstream eps g1 on s)
    rt    = (sigma / g1) coef G                   (dE/dJ = sigma q, divided by g1)
    xyz'  = (1 + eps) xyz + rt
    Xt_l  = M_l (W_l Xt_{l-1} + (1 + eps) b_l)    the forward chain RECOMPUTED on (xyz', (1+eps) latent/bias) with
                                                  the stored masks = (1 + eps) h_l + (tangent of h_l along rt)
    dW_l  = sum_rows dY_l Xt_{l-1}^T              one wgrad over the perturbed activations
    db_l  = sum_rows (1 + eps) dY_l               dlatent = sum_rows (1 + eps) (dY_0 W0lat + dY_4 W4lat)

`fused_eikonal_grads` below follows exactly these steps; tests/test_oracle_cpu.py checks it against autograd's
create_graph=True result, which pins the algebra the kernels implement before any GPU is involved.
"""
from __future__ import annotations

import torch


def fused_eikonal_grads(W, b, lat_rows, xyz, gt, delta, lam, n_total):
    """W[l] [out,in], b[l] [out] for l = 0..8 (reference shapes, skip concat at layer 4), lat_rows [M,256],
    xyz [M,3], gt [M,1] (already clamped).  Returns dict(dW, db, dlat_rows, sdf_loss, eik_loss, G)."""
    M = xyz.shape[0]
    inp = torch.cat([lat_rows, xyz], 1)
    # ---- F: forward, masks
    h, masks, X = inp, [], []
    for l in range(8):
        if l == 4:
            h = torch.cat([h, inp], 1)
        X.append(h)
        a = h @ W[l].T + b[l]
        masks.append(a > 0)
        h = torch.relu(a)
    s = h @ W[8].T + b[8]
    t = torch.tanh(s)
    inside = ((t >= -delta) & (t <= delta)).float()
    sigma = inside * (1 - t * t)
    pc = t.clamp(-delta, delta)
    sgn = torch.sign(pc - gt)
    g1 = sigma * sgn / n_total
    sdf_loss = (pc - gt).abs().sum() / n_total
    # ---- B: dgrad chain with upstream g1
    dY = [None] * 9
    dY[8] = g1
    d = g1 @ W[8]                                   # [M,512]
    jt = torch.zeros(M, 3)
    for l in range(7, -1, -1):
        d = d * masks[l]
        dY[l] = d
        if l == 4:
            jt = jt + d @ W[4][:, 509:512]
            d = d @ W[4][:, :253]
        elif l == 0:
            jt = jt + d @ W[0][:, 256:259]
        else:
            d = d @ W[l]
    # ---- per-row Eikonal scalars
    ok = (g1 != 0)
    ratio = torch.where(ok, sigma / torch.where(ok, g1, torch.ones_like(g1)), torch.zeros_like(g1))   # = N sign
    G = jt * ratio
    nG = G.norm(dim=1, keepdim=True)
    eik_loss = lam * ((1 - nG) ** 2).sum() / n_total
    pos = ok & (nG > 0)
    coef = torch.where(pos, 2 * lam / n_total * (1 - 1 / torch.where(pos, nG, torch.ones_like(nG))), torch.zeros_like(nG))
    eps = torch.where(pos, -2 * t * coef * nG * nG / torch.where(ok, g1, torch.ones_like(g1)), torch.zeros_like(nG))
    rt = ratio * coef * G
    xyz_p = (1 + eps) * xyz + rt
    rho = 1 + eps
    # ---- T: forward recomputed on the perturbed input with the stored masks
    inp_p = torch.cat([rho * lat_rows, xyz_p], 1)
    h, Xt = inp_p, []
    for l in range(8):
        if l == 4:
            h = torch.cat([h, inp_p], 1)
        Xt.append(h)
        h = (h @ W[l].T + rho * b[l]) * masks[l]
    Xt.append(h)
    # ---- wgrad over the perturbed activations, rho-weighted bias / latent sums
    dW = [dY[l].T @ Xt[l] for l in range(9)]
    db = [(rho * dY[l]).sum(0) for l in range(9)]
    dlat = (rho * dY[0]) @ W[0][:, :256] + (rho * dY[4]) @ W[4][:, 253:509]
    # the latent columns of dW0/dW4 see rho * latent (no tangent: the latent does not depend on xyz) -- already in Xt
    return {"dW": dW, "db": db, "dlat_rows": dlat, "sdf_loss": sdf_loss, "eik_loss": eik_loss, "G": G}
