"""Start / stop a single-host cluster: one controller + one engine process per GPU (or N CPU engines).

Replaces ``startCluster.sh`` (salloc + ipcontroller + ``sleep 30`` + ``srun ipengine``) and the job-submitting half of
``ipcluster_magics.py``.  There is no Slurm here: engines are local subprocesses pinned to a GPU through
``CUDA_VISIBLE_DEVICES``; readiness is a handshake (connection file + engine registration), not a sleep.

CLI:  python -m cori_intml_examples_b200.cluster start -n 8 [--cluster-id ID] [--slots K] [--cpu]
      python -m cori_intml_examples_b200.cluster stop  [--cluster-id ID]
"""
import os
import subprocess
import sys
import threading
import time

from . import protocol as P


def _gpu_count():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


class LocalCluster(object):
    """Owns the controller/engine subprocesses (or in-process thread engines for tests)."""

    def __init__(self, n_engines=None, cluster_id=None, slots=1, gpus=None, mode="process", env=None, log_dir=None,
                 isolate_gpus=True, respawn=False):
        # respawn=True: a monitor thread restarts an engine process that died (elastic recovery; the controller has already
        # re-queued or failed its tasks, the replacement registers under a new engine id and takes queued work).
        self.respawn = respawn
        self.respawns = 0
        self._engine_procs = {}          # launch slot -> Popen
        self._stopping = threading.Event()
        # isolate_gpus=True: each engine sees only its own GPU (independent HPO trials).  False: all GPUs stay visible and
        # the engine merely selects its own -- required when one data-parallel job spans the engines (DistTrain_*.ipynb).
        self.isolate_gpus = isolate_gpus
        self.cluster_id = cluster_id or "cib_%d_%d" % (os.getpid(), int(time.time() * 1000) % 100000)
        ngpu = _gpu_count()
        if gpus is None:
            gpus = list(range(ngpu)) if ngpu else []
        self.gpus = gpus
        self.n = n_engines if n_engines is not None else (len(gpus) or 2)
        self.slots, self.mode, self.env = slots, mode, dict(env or {})
        self.log_dir = log_dir or P.cluster_dir()
        self.procs, self.threads, self.engines = [], [], []
        self.controller = None

    def start(self, timeout=60):
        path = P.connection_file(self.cluster_id)
        if os.path.exists(path):
            os.remove(path)
        if self.mode == "thread":
            from .controller import Controller
            from .engine import Engine
            self.controller = Controller(self.cluster_id)
            t = threading.Thread(target=self.controller.serve, daemon=True, name="cib-controller"); t.start(); self.threads.append(t)
            P.read_connection_file(self.cluster_id, timeout)
            for i in range(self.n):
                e = Engine(self.cluster_id, gpu=None, slots=self.slots, in_process=True)
                self.engines.append(e)
                t = threading.Thread(target=e.serve, daemon=True, name="cib-engine-%d" % i); t.start(); self.threads.append(t)
        else:
            env = dict(os.environ); env.update(self.env)
            root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
            env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
            clog = open(os.path.join(self.log_dir, "controller-%s.log" % self.cluster_id), "ab")
            self.procs.append(subprocess.Popen([sys.executable, "-m", "cori_intml_examples_b200.cluster.controller", "--cluster-id", self.cluster_id],
                                               env=env, stdout=clog, stderr=subprocess.STDOUT))
            P.read_connection_file(self.cluster_id, timeout)
            self._base_env = env
            for i in range(self.n):
                self._spawn_engine(i)
            if self.respawn:
                t = threading.Thread(target=self._monitor, daemon=True, name="cib-respawn"); t.start(); self.threads.append(t)
        return self

    def _spawn_engine(self, i):
        eenv = dict(self._base_env)
        cmd = [sys.executable, "-m", "cori_intml_examples_b200.cluster.engine", "--cluster-id", self.cluster_id, "--slots", str(self.slots)]
        if self.gpus:
            g = self.gpus[i % len(self.gpus)]
            if self.isolate_gpus:
                eenv["CUDA_VISIBLE_DEVICES"] = str(g)
                cmd += ["--gpu", str(g)]
            else:
                cmd += ["--gpu", str(g), "--shared-visibility"]
        else:
            eenv["CUDA_VISIBLE_DEVICES"] = ""
        elog = open(os.path.join(self.log_dir, "engine-%s-%d.log" % (self.cluster_id, i)), "ab")
        pr = subprocess.Popen(cmd, env=eenv, stdout=elog, stderr=subprocess.STDOUT)
        self._engine_procs[i] = pr
        self.procs.append(pr)
        return pr

    def _monitor(self, period=0.5):
        while not self._stopping.wait(period):
            for i, pr in list(self._engine_procs.items()):
                if pr.poll() is not None and not self._stopping.is_set():
                    self.respawns += 1
                    self._spawn_engine(i)

    def client(self, timeout=60):
        from .client import Client
        return Client(cluster_id=self.cluster_id, timeout=timeout).wait_for_engines(self.n, timeout)

    def kill_engine(self, engine_id, sig=9):
        """Fault injection: kill the process of the engine with this *engine id* without any goodbye.
        (Engine ids are handed out in registration order, which need not be launch order.)"""
        if self.mode == "thread":
            for e in self.engines:
                if e.id == engine_id:
                    e.running = False
            return
        from .client import Client
        c = Client(cluster_id=self.cluster_id, timeout=5)
        try:
            pid = c.engine_info()[engine_id]["pid"]
        finally:
            c.close()
        import os
        os.kill(pid, sig)

    def stop(self):
        self._stopping.set()
        try:
            from .client import Client
            Client(cluster_id=self.cluster_id, timeout=2).shutdown()
        except Exception:
            pass
        deadline = time.time() + 3
        for p in self.procs:
            try:
                p.wait(max(0.05, deadline - time.time()))
            except subprocess.TimeoutExpired:
                p.kill()
        for e in self.engines:
            e.running = False
        if self.controller is not None:
            self.controller.running = False
        for t in self.threads:
            t.join(1.0)
        self.procs, self.threads, self.engines = [], [], []

    def __enter__(self):
        return self.start()

    def __exit__(self, *a):
        self.stop()


def start_cluster(n_engines=None, cluster_id=None, slots=1, gpus=None, mode="process", timeout=60):
    return LocalCluster(n_engines, cluster_id, slots, gpus, mode).start(timeout)


def stop_cluster(cluster_id=None):
    from .client import Client
    Client(cluster_id=cluster_id or P.latest_cluster_id(), timeout=5).shutdown()


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser(prog="python -m cori_intml_examples_b200.cluster")
    sub = ap.add_subparsers(dest="cmd", required=True)
    s = sub.add_parser("start"); s.add_argument("-n", "--engines", type=int, default=None); s.add_argument("--cluster-id")
    s.add_argument("--slots", type=int, default=1); s.add_argument("--cpu", action="store_true"); s.add_argument("--foreground", action="store_true")
    t = sub.add_parser("stop"); t.add_argument("--cluster-id")
    q = sub.add_parser("status"); q.add_argument("--cluster-id")
    a = ap.parse_args(argv)
    if a.cmd == "start":
        c = LocalCluster(a.engines, a.cluster_id, a.slots, gpus=[] if a.cpu else None).start()
        cl = c.client()
        print("cluster %s: %d engines registered (connection file %s)" % (c.cluster_id, len(cl.ids), P.connection_file(c.cluster_id)))
        cl.close()
        if a.foreground:
            try:
                while all(p.poll() is None for p in c.procs):
                    time.sleep(0.5)
            except KeyboardInterrupt:
                pass
            c.stop()
    elif a.cmd == "stop":
        stop_cluster(a.cluster_id)
        print("stopped")
    else:
        from .client import Client
        cl = Client(cluster_id=a.cluster_id, timeout=5)
        print(cl.queue_status())
        cl.close()
