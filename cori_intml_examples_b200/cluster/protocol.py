"""Wire protocol and connection files of the control plane.

Every ZeroMQ message is two frames: an HMAC-SHA256 signature and a pickled dict (``{"op": ..., ...}``); task
payloads (functions, arguments, results) are cloudpickle blobs *inside* that dict so that interactively defined
functions travel by value (the reference relies on that: all imports live inside ``build_and_train``,
DistHPO_mnist.ipynb:173-175).  The cluster is single-host (one 8xB200 box), so sockets are loopback TCP.

Connection file (JSON, mode 0600): ``$CIB_CLUSTER_DIR/<cluster_id>.json`` with the two URLs and the key -- the
counterpart of ipcontroller-<id>-{client,engine}.json (reference startCluster.sh:14, ipcluster_magics.py:75-79).
"""
import hashlib
import hmac
import json
import os
import pickle
import time

import cloudpickle


def cluster_dir():
    d = os.environ.get("CIB_CLUSTER_DIR") or os.path.join(os.environ.get("SCRATCH", os.path.expanduser("~")), ".cib200", "clusters")
    os.makedirs(d, exist_ok=True)
    return d


def connection_file(cluster_id):
    return os.path.join(cluster_dir(), "%s.json" % cluster_id)


def write_connection_file(cluster_id, info):
    path = connection_file(cluster_id)
    tmp = path + ".tmp"
    with open(os.open(tmp, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o600), "w") as f:
        json.dump(info, f)
    os.replace(tmp, path)
    return path


def read_connection_file(cluster_id, timeout=0.0):
    """Waits up to ``timeout`` seconds for the controller to publish its connection file (readiness handshake that
    replaces the reference's fixed ``sleep 30``, startCluster.sh:15)."""
    path = connection_file(cluster_id)
    deadline = time.time() + timeout
    while True:
        if os.path.exists(path):
            try:
                with open(path) as f:
                    return json.load(f)
            except ValueError:
                pass
        if time.time() >= deadline:
            raise TimeoutError("no connection file for cluster %r at %s" % (cluster_id, path))
        time.sleep(0.05)


def latest_cluster_id():
    files = [f for f in os.listdir(cluster_dir()) if f.endswith(".json")]
    if not files:
        raise RuntimeError("no running cluster found in %s" % cluster_dir())
    files.sort(key=lambda f: os.path.getmtime(os.path.join(cluster_dir(), f)))
    return files[-1][:-5]


def pack(msg, key):
    body = pickle.dumps(msg, protocol=pickle.HIGHEST_PROTOCOL)
    sig = hmac.new(key, body, hashlib.sha256).digest()
    return [sig, body]


def unpack(frames, key):
    sig, body = frames[-2], frames[-1]
    good = hmac.new(key, body, hashlib.sha256).digest()
    if not hmac.compare_digest(sig, good):
        raise ValueError("bad message signature")
    return pickle.loads(body)


_by_value_done = set()


def _ship_module_by_value(fn):
    """Functions from user modules that engines cannot import (scripts, notebooks, test files) travel by value, like
    interactively defined ones; library code (stdlib, site-packages, this package) stays by reference."""
    import functools
    import sys
    import sysconfig
    while isinstance(fn, functools.partial):
        fn = fn.func
    name = getattr(fn, "__module__", None)
    if not name or name == "__main__" or name in _by_value_done or name.split(".")[0] == "cori_intml_examples_b200":
        return
    mod = sys.modules.get(name)
    path = getattr(mod, "__file__", None)
    if mod is None or not path:
        return
    std = sysconfig.get_paths()["stdlib"]
    if "site-packages" in path or "dist-packages" in path or path.startswith(std):
        return
    try:
        cloudpickle.register_pickle_by_value(mod)
        _by_value_done.add(name)
    except Exception:
        pass


def dumps(obj):
    if isinstance(obj, tuple) and obj and callable(obj[0]):
        _ship_module_by_value(obj[0])
    return cloudpickle.dumps(obj)


def loads(blob):
    return pickle.loads(blob)


class RemoteError(Exception):
    """Exception raised on an engine, re-raised at ``AsyncResult.get()`` with the remote traceback attached."""

    def __init__(self, ename="RemoteError", evalue="", traceback_text="", engine_id=None):
        Exception.__init__(self, "%s: %s" % (ename, evalue))
        self.ename, self.evalue, self.traceback, self.engine_id = ename, evalue, traceback_text, engine_id

    def __str__(self):
        return "%s: %s (engine %s)\n%s" % (self.ename, self.evalue, self.engine_id, self.traceback)


class EngineError(RemoteError):
    """The engine running the task died (heartbeat lost)."""


class TaskAborted(RemoteError):
    """The task was aborted by the client (Stop button / AsyncResult.abort())."""
