"""CPU oracle for the split-NN hot path.  TEST INFRASTRUCTURE, NOT PRODUCT.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py` (its `cpu_baseline` leg and
`--impl reference`) may import this package, and only as the checker / the timed CPU baseline.
Nothing under `splitlearning_b200/` imports it; the product path fails loudly without its CUDA
library.

What it restates: the arithmetic of `/root/reference` lives in PyTorch-CPU calls
(`models.py`, `nn.CrossEntropyLoss`, `torch.optim.SGD` via `_FunctionalSGD`, `torch.optim.Adam`).
PyTorch is a third-party dependency of the reference with no version pin (no requirements file);
torch 2.11.0 CPU is what both this container and the GPU box carry.

  torch_ref.py  the same torch-CPU operator calls the reference makes, as functions with
                explicit dropout masks and explicit optimizer state (one state per Alice);
  numpy_ref.py  the published formulas of those operators in plain numpy/float64-free fp32 loops,
                for small cases; cross-checked against torch_ref.py in tests;
  protocol.py   single-process restatement of who-calls-whom (vanilla / U-shape / SISA), with
                one RNG stream per logical process;
  partition.py  LDA partition + train/test split digests (integer-exact known answers);
  gen_golden.py run HERE against `/root/reference` to write `tests/golden/*`;
  fetch_ref.sh  copies the unmodified reference into oracle/_ref/ (git-ignored; travels to the GPU box);
  ref_runner.py runs that copy - its alice/bob classes under torch.distributed.rpc, its SISA
                `bob.train_and_backward` - as the timed CPU baseline (`bench.py --impl reference`);
  cpu_baseline.py the single-process torch-CPU port, the labelled second figure / fallback.

Parity pinning: the reference has no tests and no golden vectors.  The oracle is
pinned by fixtures generated from the reference itself in this container (`gen_golden.py`:
module-level vectors from the reference's `models.py`, and recorded trajectories of the
reference's RPC entity code), committed under `tests/golden/`.
"""
