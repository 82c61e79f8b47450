"""Genetic hyper-parameter optimisation over a command-line trainer -- ``Params`` / ``Evaluator`` / ``GeneticOptimizer``.

Same user-facing surface as the Cray HPO calls in the reference (CrayHPO_rpv.ipynb:99-107,145-168,189;
CrayHPO_mnist.ipynb:41-45,62-64,75): the optimiser is a black box over a CLI -- it appends ``--flag value`` for
every parameter to ``cmd``, runs it, parses ``FoM: <float>`` from stdout (train_rpv.py:76-79) and minimises it.
Islands ("demes") of ``pop_size`` individuals evolve with tournament selection, uniform crossover, mutation and
ring migration.  Logs match the reference's whitespace-delimited files: a summary log with columns
``generation epoch best_fom avg_fom checkpoint_in checkpoint_out <params>`` and per-deme logs ``Deme<i>_<log_fn>``
with ``generation tag fitness FoM <params>`` (CrayHPO_rpv.ipynb:322,381-383).

Resource model: the reference sliced a 32-node Slurm allocation into ``nodes // nodes_per_eval`` concurrent 4-node
Horovod jobs.  Here ``nodes`` GPUs are sliced into groups of ``nodes_per_eval`` GPUs; every evaluation runs as one
process per GPU of its group (``torch.distributed.run`` when the group is larger than one) with
``CUDA_VISIBLE_DEVICES`` restricted to the group -- task parallelism over data-parallel groups.
"""
import os
import random
import re
import shlex
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

_FOM_RE = re.compile(r"FoM:\s*([-+0-9.eE]+|nan|inf)")


class Params(object):
    """``Params([[flag, default, (lo, hi) | [choices...]], ...])``; an int/float tuple is a range, a list is categorical."""

    def __init__(self, spec):
        self.names, self.defaults, self.domains = [], [], []
        for flag, default, dom in spec:
            self.names.append(flag); self.defaults.append(default)
            if isinstance(dom, tuple) and len(dom) == 2:
                kind = "int" if all(isinstance(v, int) and not isinstance(v, bool) for v in (default,) + dom) else "float"
                self.domains.append((kind, dom[0], dom[1]))
            else:
                self.domains.append(("choice", list(dom)))

    def __len__(self):
        return len(self.names)

    def sample(self, rng, i):
        d = self.domains[i]
        if d[0] == "int":
            return rng.randint(d[1], d[2])
        if d[0] == "float":
            return rng.uniform(d[1], d[2])
        return rng.choice(d[1])

    def mutate(self, rng, i, v):
        d = self.domains[i]
        if d[0] == "int":
            span = max(1, int(round(0.2 * (d[2] - d[1]))))
            return min(d[2], max(d[1], v + rng.randint(-span, span)))
        if d[0] == "float":
            return min(d[2], max(d[1], v + rng.gauss(0, 0.15 * (d[2] - d[1]))))
        return rng.choice(d[1])

    def cli(self, values):
        return " ".join("%s %s" % (n, v) for n, v in zip(self.names, values))

    def columns(self):
        return list(self.names)


def _run_cli_in_engine(script, argv):
    """Runs ``script``'s ``main(argv)`` inside a long-lived cluster engine and returns what it printed (module loaded once per engine)."""
    import contextlib
    import importlib.util
    import io
    name = "_cib_cli_" + os.path.basename(script).replace(".", "_")
    mod = sys.modules.get(name)
    if mod is None:
        spec = importlib.util.spec_from_file_location(name, script)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        mod.main(list(argv))
    return buf.getvalue()


class Evaluator(object):
    """Runs ``cmd --flag value ...`` on a slice of the GPUs and returns the parsed figure of merit.

    launcher: ``local`` / ``wlm`` / ``torchrun`` start one (torchrun) job per evaluation, as the reference's crayai evaluator does through
    the workload manager (CrayHPO_rpv.ipynb:145-151).  ``cluster`` dispatches to the LONG-LIVED engines of ``client`` instead: the
    engines are partitioned once into data-parallel groups of ``nodes_per_eval`` (each group forms a persistent communicator with
    ``hvd.init(persistent=True)``), and an evaluation is one ``main(argv)`` call per engine of a free group -- no process start, no CUDA
    context creation, no dataset re-generation per evaluation (sub-communicators inside the cluster)."""

    def __init__(self, cmd, nodes=1, nodes_per_eval=1, alloc_args="", launcher="local", verbose=False, timeout=None,
                 gpus=None, env=None, run_dir=None, client=None):
        self.client, self._groups = client, None
        if launcher == "cluster" and client is None:
            raise ValueError("launcher='cluster' needs client=<cluster.Client> with at least nodes engines")
        self.cmd, self.nodes, self.nodes_per_eval = cmd, int(nodes), int(nodes_per_eval)
        self.alloc_args, self.launcher, self.verbose, self.timeout = alloc_args, launcher, verbose, timeout
        self.n_concurrent = max(1, self.nodes // self.nodes_per_eval)
        self.gpus = list(gpus) if gpus is not None else list(range(self.nodes))
        self.env = dict(env or {})
        self.run_dir = run_dir
        self._slots = list(range(self.n_concurrent))
        self._lock = threading.Lock()
        self._port = 29600 + (os.getpid() % 200) * 16
        self.n_evals = 0

    def _acquire(self):
        while True:
            with self._lock:
                if self._slots:
                    return self._slots.pop(0)
            time.sleep(0.01)

    def _release(self, slot):
        with self._lock:
            self._slots.append(slot)

    def command(self, params, values, slot):
        base = self.cmd + " " + params.cli(values)
        if self.nodes_per_eval > 1 and self.launcher in ("wlm", "torchrun", "local"):
            exe = base.split()[0]
            rest = base[len(exe):].strip() if exe in ("python", "python3", sys.executable) else base
            return "%s -m torch.distributed.run --nnodes=1 --nproc-per-node %d --master-addr 127.0.0.1 --master-port %d %s" % (
                sys.executable, self.nodes_per_eval, self._port + slot, rest)
        if base.startswith("python "):
            base = sys.executable + base[6:]
        return base

    # ---------------------------------------------------------------- cluster launcher
    def _cluster_groups(self):
        """Partition the engines into data-parallel groups once; every group gets its own persistent communicator."""
        with self._lock:
            if self._groups is not None:
                return self._groups
            ids = list(self.client.ids)[:self.n_concurrent * self.nodes_per_eval]
            if len(ids) < self.n_concurrent * self.nodes_per_eval:
                raise RuntimeError("cluster has %d engines, %d needed" % (len(ids), self.n_concurrent * self.nodes_per_eval))
            groups = [ids[g * self.nodes_per_eval:(g + 1) * self.nodes_per_eval] for g in range(self.n_concurrent)]
            for grp in groups:
                view = self.client[grp]
                if self.nodes_per_eval > 1:
                    view.setup_dist()
                view.execute("from cori_intml_examples_b200.parallel import dist as _cib_hvd\n_cib_hvd.init(persistent=True)", block=True, silent=True)
            self._groups = groups
            return groups

    def _evaluate_on_cluster(self, params, values, slot):
        tokens = shlex.split(self.cmd + " " + params.cli(values))
        if tokens and os.path.basename(tokens[0]).startswith("python"):
            tokens = tokens[1:]
        script, argv = tokens[0], tokens[1:]
        view = self.client[self._cluster_groups()[slot]]
        try:
            outs = view.apply_async(_run_cli_in_engine, script, argv).get(self.timeout)
        except Exception as e:                      # a failed training (bad hyper-parameters, remote error) is an infinitely bad individual
            if self.verbose:
                print("[evaluator group %d] evaluation failed: %s" % (slot, e))
            return ""
        return "\n".join(o for o in (outs if isinstance(outs, (list, tuple)) else [outs]) if isinstance(o, str))

    def evaluate(self, params, values):
        slot = self._acquire()
        try:
            if self.launcher == "cluster":
                t0 = time.time()
                out = self._evaluate_on_cluster(params, values, slot)
                m = _FOM_RE.findall(out)
                fom = float("inf")
                for tok in reversed(m):
                    try:
                        fom = float(tok); break
                    except ValueError:
                        continue
                if fom != fom:
                    fom = float("inf")
                with self._lock:
                    self.n_evals += 1
                return fom, time.time() - t0
            group = self.gpus[slot * self.nodes_per_eval:(slot + 1) * self.nodes_per_eval]
            env = dict(os.environ); env.update(self.env)
            if group and env.get("CIB_HPO_CPU") != "1":
                env["CUDA_VISIBLE_DEVICES"] = ",".join(str(g) for g in group)
            cmd = self.command(params, values, slot)
            if self.verbose:
                print("[evaluator slot %d] %s" % (slot, cmd))
            t0 = time.time()
            try:
                pr = subprocess.run(shlex.split(cmd), capture_output=True, text=True, timeout=self.timeout, env=env, cwd=self.run_dir)
                out = pr.stdout
            except subprocess.TimeoutExpired as e:
                out = e.stdout or ""
            m = _FOM_RE.findall(out if isinstance(out, str) else out.decode())
            fom = float("inf")
            for tok in reversed(m):                    # last well-formed value wins (ranks may interleave their output)
                try:
                    fom = float(tok); break
                except ValueError:
                    continue
            if fom != fom:
                fom = float("inf")
            with self._lock:
                self.n_evals += 1
            return fom, time.time() - t0
        finally:
            self._release(slot)


class GeneticOptimizer(object):
    def __init__(self, evaluator, pop_size=8, num_demes=1, generations=4, mutation_rate=0.05, crossover_rate=0.33,
                 verbose=False, log_fn=None, seed=0, migration=True):
        self.evaluator, self.pop_size, self.num_demes, self.generations = evaluator, int(pop_size), int(num_demes), int(generations)
        self.mutation_rate, self.crossover_rate, self.verbose, self.log_fn = float(mutation_rate), float(crossover_rate), verbose, log_fn
        self.rng = random.Random(seed)
        self.migration = migration
        self.history = []           # per generation: dict(best_fom, avg_fom, best_params)
        self.evaluations = []       # every individual: dict(generation, deme, tag, fom, params, secs)
        self.best_fom, self.best_params = float("inf"), None

    # ---------------------------------------------------------------- GA operators
    def _child(self, params, pop, foms):
        def pick():
            a, b = self.rng.randrange(len(pop)), self.rng.randrange(len(pop))
            return pop[a] if foms[a] <= foms[b] else pop[b]
        p1, p2 = pick(), pick()
        child = [p2[i] if self.rng.random() < self.crossover_rate else p1[i] for i in range(len(params))]
        for i in range(len(params)):
            if self.rng.random() < self.mutation_rate:
                child[i] = params.mutate(self.rng, i, child[i])
        return child

    def _fmt(self, v):
        return ("%.6g" % v) if isinstance(v, float) else str(v)

    def optimize(self, params):
        P, D = self.pop_size, self.num_demes
        demes = []
        for d in range(D):
            pop = [[params.sample(self.rng, i) for i in range(len(params))] for _ in range(P)]
            if d == 0:
                pop[0] = list(params.defaults)
            demes.append(pop)
        summary = deme_logs = None
        if self.log_fn:
            summary = open(self.log_fn, "w")
            summary.write(" ".join(["generation", "epoch", "best_fom", "avg_fom", "checkpoint_in", "checkpoint_out"] + params.columns()) + "\n")
            ddir, base = os.path.split(self.log_fn)
            deme_logs = [open(os.path.join(ddir, "Deme%d_%s" % (d + 1, base)), "w") for d in range(D)]
            for f in deme_logs:
                f.write(" ".join(["generation", "tag", "fitness", "FoM"] + params.columns()) + "\n")
        pool = ThreadPoolExecutor(max_workers=self.evaluator.n_concurrent)
        try:
            for gen in range(self.generations):
                jobs = [(d, k, pool.submit(self.evaluator.evaluate, params, ind)) for d, pop in enumerate(demes) for k, ind in enumerate(pop)]
                foms = [[None] * P for _ in range(D)]
                for d, k, fut in jobs:
                    fom, secs = fut.result()
                    foms[d][k] = fom
                    self.evaluations.append({"generation": gen, "deme": d, "tag": k, "fom": fom, "params": list(demes[d][k]), "secs": secs})
                    if deme_logs:
                        fit = 1.0 / fom if fom not in (0.0, float("inf")) else 0.0
                        deme_logs[d].write(" ".join([str(gen), str(k), "%.6g" % fit, "%.6g" % fom] + [self._fmt(v) for v in demes[d][k]]) + "\n")
                        deme_logs[d].flush()
                    if fom < self.best_fom:
                        self.best_fom, self.best_params = fom, list(demes[d][k])
                flat = [f for row in foms for f in row]
                finite = [f for f in flat if f != float("inf")]
                avg = sum(finite) / len(finite) if finite else float("inf")
                gbest = min(flat)
                gd, gk = next((d, k) for d in range(D) for k in range(P) if foms[d][k] == gbest)
                self.history.append({"generation": gen, "best_fom": gbest, "avg_fom": avg, "best_params": list(demes[gd][gk])})
                if summary:
                    summary.write(" ".join([str(gen), str(gen), "%.6g" % gbest, "%.6g" % avg, "NaN", "NaN"] + [self._fmt(v) for v in demes[gd][gk]]) + "\n")
                    summary.flush()
                if self.verbose:
                    print("generation %d: best_fom %.6g avg_fom %.6g (%d evaluations so far)" % (gen, gbest, avg, len(self.evaluations)))
                if gen == self.generations - 1:
                    break
                # next generation: elitism + children, then ring migration of each deme's best
                new_demes = []
                for d in range(D):
                    order = sorted(range(P), key=lambda k: foms[d][k])
                    nxt = [list(demes[d][order[0]])]
                    while len(nxt) < P:
                        nxt.append(self._child(params, demes[d], foms[d]))
                    new_demes.append(nxt)
                if self.migration and D > 1:
                    bests = [list(demes[d][min(range(P), key=lambda k: foms[d][k])]) for d in range(D)]
                    for d in range(D):
                        new_demes[(d + 1) % D][-1] = bests[d]
                demes = new_demes
        finally:
            pool.shutdown(wait=True)
            if summary:
                summary.close()
                for f in deme_logs:
                    f.close()
        return dict(zip(params.names, self.best_params)) if self.best_params is not None else None
