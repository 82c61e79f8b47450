#!/usr/bin/env python
"""Genetic HPO over the MNIST command-line trainer -- CrayHPO_mnist.ipynb: Params with --h1/--h2/--h3/--dropout/--optimizer
(:41-45), Evaluator('python train.py --epochs 4', nodes=32) (:58,75) -- the reference never shipped that train.py; ours is
train.py / train_mnist.py at the repository root -- and GeneticOptimizer(pop_size=4, num_demes=4, generations=16) (:62-64)."""
import os

import _common


def main(argv=None):
    ap = _common.parser(__doc__)
    ap.add_argument("--nodes", type=int, default=None, help="GPUs to spread evaluations over (default: all)")
    ap.add_argument("--generations", type=int, default=16)
    ap.add_argument("--pop-size", type=int, default=4)
    ap.add_argument("--num-demes", type=int, default=4)
    a = ap.parse_args(argv)
    from cori_intml_examples_b200.hpo import genetic
    dev = _common.resolve_device(a.device)
    nodes = a.nodes or (8 if dev == "gpu" else 2)
    params = genetic.Params([["--h1", 4, (2, 64)], ["--h2", 8, (2, 64)], ["--h3", 32, (8, 128)], ["--dropout", 0.5, (0.0, 0.9)],
                             ["--optimizer", "Adadelta", ["Adadelta", "Adam", "Nadam"]]])
    cmd = "python %s --epochs 4 --fom last --verbose 0 --device %s" % (os.path.join(_common.ROOT, "train.py"), dev)
    if a.quick:
        a.generations, a.pop_size, a.num_demes = 2, 2, 1
        params = genetic.Params([["--h1", 4, (2, 8)], ["--h3", 8, (8, 16)], ["--optimizer", "Adam", ["Adam", "Nadam"]]])
        cmd = "python %s --epochs 1 --n-train 128 --n-test 16 --batch-size 32 --verbose 0 --device %s" % (os.path.join(_common.ROOT, "train.py"), dev)
    env = {"CIB_HPO_CPU": "1", "CUDA_VISIBLE_DEVICES": ""} if dev == "cpu" else {}
    ev = genetic.Evaluator(cmd, nodes=nodes, nodes_per_eval=1, launcher="local", verbose=True, env=env)
    log = os.path.join(a.out_dir, "genetic_mnist_%d.log" % os.getpid())
    opt = genetic.GeneticOptimizer(ev, pop_size=a.pop_size, num_demes=a.num_demes, generations=a.generations, mutation_rate=0.05,
                                   crossover_rate=0.33, verbose=True, log_fn=log)
    opt.optimize(params)
    print("Best FoM %.6f with %s" % (opt.best_fom, opt.best_params))
    print("logs:", log)
    return opt


if __name__ == "__main__":
    main()
