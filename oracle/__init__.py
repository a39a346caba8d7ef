"""CPU oracle for the DDP toy-model step -- TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
leg may import this package.  See ddp_oracle.c for the restated algorithms and
torch_ref.py for the stock-PyTorch restatement of the reference loop.
"""
from .oracle import Oracle, build_oracle, TOY_DIMS  # noqa: F401
