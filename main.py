"""Entry point: ``python main.py --params utils/cifar_params.yaml``  (reference main.py:84-92).

With more than one visible GPU (or ``n_gpus: N`` in the YAML) the script re-launches itself as one process per GPU
(``torch.distributed`` over NCCL for rendezvous; the aggregation hot paths use the fused peer-memory kernels), so the
CLI contract stays the single command of the reference.
"""
from __future__ import annotations

import argparse
import os
import sys


def main(argv=None) -> int:
    parser = argparse.ArgumentParser(description="PPDL")
    parser.add_argument("--params", dest="params", required=True)
    parser.add_argument("--rounds", type=int, default=None, help="stop after this many rounds (default: all)")
    parser.add_argument("--gpus", type=int, default=None, help="override n_gpus from the YAML")
    parser.add_argument("--quiet", action="store_true")
    args = parser.parse_args(argv)

    print("Start training")
    from dba_mod_b200.config import Params
    from dba_mod_b200.parallel.launcher import maybe_spawn, make_comm

    params = Params.from_yaml(args.params)
    if args.gpus is not None:
        params["n_gpus"] = args.gpus
    if maybe_spawn(params, sys.argv):          # parent of a multi-GPU run: children did the work
        return 0
    comm = make_comm()
    from dba_mod_b200.fl.simulator import Simulator
    sim = Simulator(params, comm=comm, quiet=args.quiet)
    try:
        sim.run(max_rounds=args.rounds)
    finally:
        sim.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
