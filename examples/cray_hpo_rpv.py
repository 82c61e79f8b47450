#!/usr/bin/env python
"""Genetic HPO over multi-GPU RPV trainings -- CrayHPO_rpv.ipynb: 7 parameters incl. integer ranges (:99-107), Evaluator over
`python train_rpv.py --n-epochs 4 --fom best` with nodes=32, nodes_per_eval=4, launcher='wlm' (:62-63,145-151),
GeneticOptimizer(pop_size=8, num_demes=4, generations=4, mutation_rate=0.05, crossover_rate=0.33, log_fn) (:84-88,161-168),
optimize (:189), and the summary / per-deme logs read back with pandas (:322,381-383).

On one 8xB200 box a "node" is a GPU: `--nodes 8 --nodes-per-eval 4` runs two concurrent 4-GPU data-parallel evaluations (the
hybrid task-parallel-over-DP strategy of SURVEY 2.5)."""
import os

import _common


def main(argv=None):
    ap = _common.parser(__doc__)
    ap.add_argument("--nodes", type=int, default=None)
    ap.add_argument("--nodes-per-eval", type=int, default=None)
    ap.add_argument("--generations", type=int, default=4)
    ap.add_argument("--pop-size", type=int, default=8)
    ap.add_argument("--num-demes", type=int, default=4)
    ap.add_argument("--launcher", default="wlm", choices=["wlm", "cluster"],
                    help="wlm: one torchrun job per evaluation (the reference's way); cluster: long-lived engines in persistent DP groups")
    ap.add_argument("--channels", type=int, default=None, help="channels of the synthetic images (BASELINE.json: 3)")
    a = ap.parse_args(argv)
    from cori_intml_examples_b200.hpo import genetic
    dev = _common.resolve_device(a.device)
    nodes = a.nodes or (8 if dev == "gpu" else 2)
    per = a.nodes_per_eval or (4 if dev == "gpu" and nodes >= 4 else 1)
    script = os.path.join(_common.ROOT, "train_rpv.py")
    params = genetic.Params([["--h1", 16, (4, 64)], ["--h2", 32, (4, 64)], ["--h3", 64, (8, 128)], ["--h4", 128, (32, 256)],
                             ["--dropout", 0.2, (0.0, 0.9)], ["--optimizer", "Adam", ["Adam", "Nadam", "Adadelta"]],
                             ["--lr", 0.001, (1e-4, 1e-2)]])
    cmd = "python %s --input-dir %s --n-epochs 4 --fom best --lr-scaling linear --verbose 0 --device %s" % (script, a.out_dir, dev)
    if a.quick:
        a.generations, a.pop_size, a.num_demes, per = 2, 2, 1, min(per, 2)
        params = genetic.Params([["--h1", 4, (3, 8)], ["--dropout", 0.2, (0.0, 0.5)], ["--optimizer", "Adam", ["Adam", "Nadam"]]])
        cmd = ("python %s --input-dir %s --n-train 64 --n-valid 32 --h2 4 --h3 4 --h4 8 --batch-size 16 --n-epochs 1 --fom best "
               "--verbose 0 --device %s" % (script, a.out_dir, dev))
    if a.channels:
        cmd += " --channels %d" % a.channels
    env = {"CIB_HPO_CPU": "1", "CUDA_VISIBLE_DEVICES": ""} if dev == "cpu" else {}
    log = os.path.join(a.out_dir, "genetic_rpv_%d.log" % os.getpid())
    import contextlib
    import time
    from cori_intml_examples_b200.cluster import LocalCluster
    with contextlib.ExitStack() as stack, _common.Recorder(a, "cray_hpo_rpv") as rec:
        client = None
        if a.launcher == "cluster":
            # one engine per "node" (GPU), all GPUs visible to every engine so that a group's ranks can map each other's memory
            cl = stack.enter_context(LocalCluster(nodes, mode="process", gpus=None if dev == "gpu" else [], isolate_gpus=False, env=env))
            client = cl.client()
        ev = genetic.Evaluator(cmd, nodes=nodes, nodes_per_eval=per, launcher=a.launcher, verbose=a.launcher != "cluster", env=env, client=client)
        opt = genetic.GeneticOptimizer(ev, pop_size=a.pop_size, num_demes=a.num_demes, generations=a.generations, mutation_rate=0.05,
                                       crossover_rate=0.33, verbose=True, log_fn=log)
        t0 = time.time()
        with rec.under_load():
            opt.optimize(params)
        wall = time.time() - t0
        secs = [e["secs"] for e in opt.evaluations]
        rec["launcher"] = a.launcher; rec["nodes"] = nodes; rec["nodes_per_eval"] = per; rec["concurrent_evaluations"] = ev.n_concurrent
        rec["evaluations"] = ev.n_evals; rec["optimize_wall_s"] = wall
        rec["evaluations_per_hour"] = ev.n_evals / wall * 3600.0
        rec["reference_evaluations_per_hour"] = 41.2          # CrayHPO_rpv.ipynb:180-181: 128 evaluations in 11173 s on 32 Haswell nodes
        rec["eval_secs_mean"] = sum(secs) / max(1, len(secs)); rec["eval_secs_min"] = min(secs); rec["eval_secs_max"] = max(secs)
        rec["best_fom"] = opt.best_fom; rec["best_params"] = opt.best_params; rec["history"] = opt.history
        rec["finite_foms"] = sum(1 for e in opt.evaluations if e["fom"] != float("inf"))
        rec["command"] = cmd
    print("Best FoM %.6f with %s  (%d evaluations in %.1f s = %.1f evaluations/hour)" % (opt.best_fom, opt.best_params, ev.n_evals, wall, ev.n_evals / wall * 3600))
    try:
        import pandas as pd
        print(pd.read_csv(log, sep=r"\s+").to_string(index=False))
        print(pd.read_csv(os.path.join(os.path.dirname(log), "Deme1_" + os.path.basename(log)), sep=r"\s+").head().to_string(index=False))
    except Exception as e:      # pandas is optional
        print("(pandas not available: %s)" % e)
    return opt


if __name__ == "__main__":
    main()
