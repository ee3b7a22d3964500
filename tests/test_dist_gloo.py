"""world_size-2 gloo test (CPU) of the scene-sharded step logic: each rank processes its own
scenes, normalises by the GLOBAL row count, and the decoder gradients are all-reduced with SUM.
The per-rank compute is the oracle (the CUDA kernels need a GPU); what is under test is the host
logic in dsdf_b200.dist: partition, global N, SUM semantics, latent rows never leaving the owner."""
import os
import socket

import torch
import torch.distributed as dist_
import torch.multiprocessing as mp

from common import cos, spec_for


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, S, B, P, out_dir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.join(here, ".."), os.path.join(here, "..", "medical-shape-disentanglement_b200")):
        sys.path.insert(0, p)
    from dsdf_b200 import dist, synth
    from oracle import train as otrain
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist_.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    spec = spec_for(True)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0)
    tr = otrain.OracleTrainer(spec, cfg, S, seed=0)              # replicas start identical
    data = synth.scene_samples(S, P, seed=1)
    mine = dist.owned_scenes(S, rank, world)
    nb = dist.local_batch(B, rank, world)
    ids = torch.tensor(list(mine)[:nb])
    n_total = dist.global_rows(nb * P)
    assert n_total == B * P
    tr.opt.zero_grad()
    tr.forward_loss(data[ids], ids, epoch=60, n_total=n_total)
    flat = torch.cat([p.grad.flatten() for p in tr.params.values()])
    dist.allreduce_decoder_grads(flat)
    lg = tr.latent_grad()
    # latent grads exist only on rows this rank owns
    owned = torch.zeros(S, dtype=torch.bool); owned[list(mine)] = True
    assert torch.all(lg[~owned] == 0)
    table = dist.gather_latents_for_checkpoint(tr.lat_vecs.weight.data[list(mine)], S, rank, world)
    torch.save({"flat": flat, "lat_grad": lg, "ids": ids, "table": table}, os.path.join(out_dir, f"r{rank}.pt"))
    dist_.destroy_process_group()


def test_sharded_step_equals_single_process(tmp_path):
    from dsdf_b200 import synth
    from oracle import train as otrain
    S, B, P, W = 8, 4, 128, 2
    mp.spawn(_worker, args=(W, _free_port(), S, B, P, str(tmp_path)), nprocs=W, join=True)
    r = [torch.load(os.path.join(tmp_path, f"r{i}.pt")) for i in range(W)]
    assert torch.equal(r[0]["flat"], r[1]["flat"])               # replicas see the same summed grad
    ids = torch.cat([r[0]["ids"], r[1]["ids"]])
    spec = spec_for(True)
    cfg = otrain.TrainConfig(samples_per_scene=P, scenes_per_batch=B, code_bound=1.0, grad_clip=1.0)
    tr = otrain.OracleTrainer(spec, cfg, S, seed=0)
    data = synth.scene_samples(S, P, seed=1)
    tr.opt.zero_grad()
    tr.forward_loss(data[ids], ids, epoch=60)
    flat = torch.cat([p.grad.flatten() for p in tr.params.values()])
    assert cos(flat, r[0]["flat"]) > 0.999999
    torch.testing.assert_close(flat, r[0]["flat"], rtol=1e-3, atol=1e-7)
    torch.testing.assert_close(tr.latent_grad(), r[0]["lat_grad"] + r[1]["lat_grad"], rtol=1e-4, atol=1e-9)
    assert torch.equal(r[0]["table"], tr.lat_vecs.weight.data) or torch.allclose(r[0]["table"], tr.lat_vecs.weight.data)
