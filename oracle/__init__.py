"""CPU/fp32 oracle for the DeepSDF decoder hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(`medical-shape-disentanglement_b200/`) imports from here.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py` (the `cpu_baseline` leg and
`--impl reference`) may import this package, and only as the checker / the timed
CPU baseline -- never as the implementation being shipped or measured.

What it restates (plain PyTorch fp32, device agnostic), each function citing the
reference file:line it follows:

  deepsdf.py      networks/deep_sdf_decoder.py:10-109   (Decoder ctor + forward)
  train.py        train_deep_sdf.py:429-454,555-616,731,780-784 (auto-decoder step)
  mesh_grid.py    deep_sdf/mesh.py:35-71, deep_sdf/utils.py:86-97 (grid + decode)
  reconstruct.py  reconstruct.py:16-151, deep_sdf/data.py:112-136 (latent fit)
  siren.py        networks/siren_decoder.py:83-237, networks/modules.py:4-14

Parity pinning: the reference ships no tests / golden vectors, so
the oracle is pinned against the *live* reference modules imported from
/root/reference (tests/test_oracle_vs_reference.py, skipped where the reference
tree is absent) and against fixtures generated from them
(tests/golden/make_golden.py -> tests/golden/*.npz, committed).
"""
