"""Client API: ``Client``, ``DirectView``, ``LoadBalancedView`` and ``AsyncResult`` handles.

Mirrors the IPyParallel surface the reference notebooks use: ``Client(timeout=60, cluster_id=...)``,
``c.ids``, ``c[i]`` / ``c[:]`` with ``.execute`` (= ``%%px``) / ``.get`` / ``.push`` / ``.pull`` / ``.apply``,
``c.load_balanced_view().apply(f, *a, **kw)`` -> ``AsyncResult`` with ``ready() done() get() wait() stdout stderr
started completed metadata data abort()`` (DistHPO_mnist.ipynb:240-272,333-344; hpo_widgets.py:260,402).
"""
import datetime
import queue
import threading
import time
import uuid

import zmq

from . import protocol as P


class AsyncResult(object):
    """Handle for one task (or, for multi-engine direct views, a list of per-engine tasks)."""

    def __init__(self, client, task_ids, targets, single):
        self._client, self._ids, self._targets, self._single = client, task_ids, targets, single
        self._tasks = [client._tasks[t] for t in task_ids]

    # ---- state
    def ready(self):
        return all(t["done"].is_set() for t in self._tasks)

    def done(self):
        return self.ready()

    def successful(self):
        if not self.ready():
            raise ValueError("result is not ready")
        return all(t["status"] == "ok" for t in self._tasks)

    def wait(self, timeout=None):
        deadline = None if timeout is None else time.time() + timeout
        for t in self._tasks:
            t["done"].wait(None if deadline is None else max(0.0, deadline - time.time()))
        return self.ready()

    def get(self, timeout=None):
        if not self.wait(timeout):
            raise TimeoutError("result not ready after %s s" % timeout)
        outs = []
        for t in self._tasks:
            if t["status"] != "ok":
                ename, evalue, tb = t["error"] or ("RemoteError", "", "")
                cls = {"EngineError": P.EngineError, "TaskAborted": P.TaskAborted}.get(ename, P.RemoteError)
                raise cls(ename, evalue, tb, t["engine_id"])
            outs.append(t["result"])
        return outs[0] if self._single else outs

    r = property(lambda self: self.get())
    result = get

    def abort(self):
        for tid, t in zip(self._ids, self._tasks):
            if not t["done"].is_set():
                self._client._send({"op": "abort", "task_id": tid})

    # ---- metadata
    def _field(self, name):
        vals = [t[name] for t in self._tasks]
        return vals[0] if self._single else vals

    @property
    def stdout(self):
        return self._field("stdout")

    @property
    def stderr(self):
        return self._field("stderr")

    @property
    def data(self):
        """Latest dict published by the task with ``publish_data`` ({} before the first publication)."""
        return self._field("data")

    @property
    def engine_id(self):
        return self._field("engine_id")

    @staticmethod
    def _dt(ts):
        return None if ts is None else datetime.datetime.fromtimestamp(ts)

    @property
    def submitted(self):
        v = [self._dt(t["submitted"]) for t in self._tasks]
        return v[0] if self._single else v

    @property
    def started(self):
        v = [self._dt(t["started"]) for t in self._tasks]
        return v[0] if self._single else v

    @property
    def completed(self):
        v = [self._dt(t["completed"]) for t in self._tasks]
        return v[0] if self._single else v

    @property
    def elapsed(self):
        ts = [t for t in self._tasks if t["started"] is not None]
        if not ts:
            return 0.0
        end = max((t["completed"] or time.time()) for t in ts)
        return end - min(t["started"] for t in ts)

    @property
    def metadata(self):
        md = [{"msg_id": tid, "engine_id": t["engine_id"], "submitted": self._dt(t["submitted"]), "started": self._dt(t["started"]),
               "completed": self._dt(t["completed"]), "status": t["status"], "stdout": t["stdout"], "stderr": t["stderr"],
               "data": t["data"], "error": t["error"], "requeues": t["requeues"]} for tid, t in zip(self._ids, self._tasks)]
        return md[0] if self._single else md

    def display_outputs(self):
        for t in self._tasks:
            if t["stdout"]:
                print("[stdout:%s] %s" % (t["engine_id"], t["stdout"]), end="" if t["stdout"].endswith("\n") else "\n")
            if t["stderr"]:
                print("[stderr:%s] %s" % (t["engine_id"], t["stderr"]), end="" if t["stderr"].endswith("\n") else "\n")

    def __repr__(self):
        return "<AsyncResult %s: %s>" % (",".join(i[:8] for i in self._ids), "finished" if self.ready() else "pending")


class View(object):
    def __init__(self, client):
        self.client = client
        self.block = False
        self.retries = 0


class DirectView(View):
    """Engines addressed explicitly; every call fans out to all targets (``c[:]``) or goes to the one (``c[i]``)."""

    def __init__(self, client, targets, single):
        View.__init__(self, client)
        self.targets, self._single = list(targets), single

    def _fan(self, kind, payload_fn, block):
        ar = self.client._submit_many(kind, [(t, payload_fn(t)) for t in self.targets], self._single)
        if block:
            ar.get()
        return ar

    def execute(self, code, block=True, silent=False):
        """Run source in each engine's persistent namespace (the ``%%px`` cell magic); echoes ``[stdout:i]`` blocks."""
        ar = self._fan("execute", lambda t: P.dumps(code), False)
        if block:
            ar.wait()
            if not silent:
                ar.display_outputs()
            ar.get()
        return ar

    def setup_dist(self, master_port=None):
        """Give the engines of this view a rank each so that a bare ``hvd.init()`` inside ``%%px`` forms one communicator
        over them -- the role ``srun`` + MPI_Init play for the reference's engines (startCluster.sh:18,
        DistTrain_rpv.ipynb:136-141).  Rank = position in the view; returns the rendezvous port."""
        if master_port is None:
            import socket
            with socket.socket() as sk:
                sk.bind(("127.0.0.1", 0))
                master_port = sk.getsockname()[1]
        n = len(self.targets)
        code = ("import os as _os\n_os.environ.update(RANK=str(%r.index(engine_id)), WORLD_SIZE='%d', MASTER_ADDR='127.0.0.1', "
                "MASTER_PORT='%d')\n" % (list(self.targets), n, master_port))
        self.execute(code, block=True, silent=True)
        return master_port

    def run(self, filename, block=True):
        with open(filename) as f:
            return self.execute(f.read(), block=block)

    def apply_async(self, f, *args, **kwargs):
        return self._fan("apply", lambda t: P.dumps((f, args, kwargs)), False)

    def apply_sync(self, f, *args, **kwargs):
        return self.apply_async(f, *args, **kwargs).get()

    def apply(self, f, *args, **kwargs):
        ar = self.apply_async(f, *args, **kwargs)
        return ar.get() if self.block else ar

    def push(self, ns, block=True):
        return self._fan("push", lambda t: P.dumps(dict(ns)), block)

    def pull(self, names, block=True):
        ar = self._fan("pull", lambda t: P.dumps(names), False)
        return ar.get() if block else ar

    def get(self, expr):
        """``c[0].get('history.epoch')`` / ``c[:].get('history.history')`` (DistTrain_rpv.ipynb:310-311)."""
        return self.pull(expr, block=True)

    def __getitem__(self, name):
        return self.pull(name, block=True)

    def __setitem__(self, name, value):
        self.push({name: value}, block=True)

    def scatter(self, name, seq, block=True):
        n = len(self.targets)
        chunks = [seq[i * len(seq) // n:(i + 1) * len(seq) // n] for i in range(n)]
        return self.client._submit_many("push", [(t, P.dumps({name: c})) for t, c in zip(self.targets, chunks)], False).wait() if block else None

    def gather(self, name, block=True):
        parts = DirectView(self.client, self.targets, False).pull(name)
        out = []
        for p in parts:
            out.extend(p)
        return out

    def map_sync(self, f, *seqs):
        return [ar.get() for ar in [self.client.load_balanced_view(self.targets).apply(f, *a) for a in zip(*seqs)]]

    def __len__(self):
        return len(self.targets)


class LoadBalancedView(View):
    """Tasks queue at the controller and go to the least-loaded engine with a free slot (DistHPO_mnist.ipynb:240-255)."""

    def __init__(self, client, targets=None):
        View.__init__(self, client)
        self.targets = targets

    def apply_async(self, f, *args, **kwargs):
        return self.client._submit_many("apply", [("lb", P.dumps((f, args, kwargs)))], True, retries=self.retries)

    def apply(self, f, *args, **kwargs):
        ar = self.apply_async(f, *args, **kwargs)
        return ar.get() if self.block else ar

    def apply_sync(self, f, *args, **kwargs):
        return self.apply_async(f, *args, **kwargs).get()

    def map_async(self, f, *seqs):
        return [self.apply_async(f, *a) for a in zip(*seqs)]

    def map(self, f, *seqs, **kw):
        ars = self.map_async(f, *seqs)
        return [a.get() for a in ars] if kw.get("block", True) else ars

    map_sync = map


class Client(object):
    def __init__(self, cluster_id=None, timeout=60, profile=None, **kwargs):
        self.cluster_id = cluster_id or P.latest_cluster_id()
        info = P.read_connection_file(self.cluster_id, timeout)
        self._key = bytes.fromhex(info["key"])
        self._ctx = zmq.Context.instance()
        self._sock = self._ctx.socket(zmq.DEALER)
        self._sock.setsockopt(zmq.LINGER, 0)
        self._sock.connect(info["client_url"])
        self._outq = queue.Queue()
        self._tasks = {}
        self._replies = {}
        self._reply_cv = threading.Condition()
        self._engines = []
        self._engine_info = {}
        self._closed = False
        self._thread = threading.Thread(target=self._recv_loop, daemon=True, name="cib-client-recv")
        self._thread.start()
        self._send({"op": "connect"})
        if self._wait_reply("connect_reply", timeout) is None:
            raise TimeoutError("controller %s did not answer within %s s" % (self.cluster_id, timeout))

    # ------------------------------------------------------------------ plumbing
    def _send(self, msg):
        # ZeroMQ sockets are not thread-safe: only the I/O thread touches the socket, everybody else queues
        self._outq.put(msg)

    def _wait_reply(self, op, timeout):
        deadline = time.time() + timeout
        with self._reply_cv:
            while op not in self._replies:
                left = deadline - time.time()
                if left <= 0:
                    return None
                self._reply_cv.wait(left)
            return self._replies.pop(op)

    def _recv_loop(self):
        poller = zmq.Poller(); poller.register(self._sock, zmq.POLLIN)
        while True:
            try:
                while True:
                    try:
                        self._sock.send_multipart(P.pack(self._outq.get_nowait(), self._key))
                    except queue.Empty:
                        break
                if self._closed:
                    break
                if not poller.poll(5):
                    continue
                frames = self._sock.recv_multipart(zmq.NOBLOCK)
                msg = P.unpack(frames, self._key)
            except zmq.Again:
                continue
            except Exception:
                if self._closed:
                    break
                continue
            op = msg["op"]
            t = self._tasks.get(msg.get("task_id"))
            if op == "result" and t is not None:
                t.update(status=msg["status"], error=msg.get("error"), engine_id=msg.get("engine_id"),
                         stdout=msg.get("stdout", ""), stderr=msg.get("stderr", ""),
                         started=msg.get("started") or t["started"], completed=msg.get("completed"))
                if msg["status"] == "ok":
                    try:
                        t["result"] = P.loads(msg["payload"])
                    except Exception as e:          # result not unpicklable on the client
                        t["status"], t["error"] = "error", (type(e).__name__, str(e), "")
                t["done"].set()
            elif op == "started" and t is not None:
                t["started"], t["engine_id"] = msg["time"], msg.get("engine_id")
            elif op == "data" and t is not None:
                try:
                    t["data"] = P.loads(msg["data"])
                except Exception:
                    pass
            elif op == "stream" and t is not None:
                t[msg["name"]] += msg["text"]
            elif op == "requeued" and t is not None:
                t["requeues"] += 1; t["started"] = None; t["stdout"] = t["stderr"] = ""
            elif op in ("connect_reply", "status_reply", "shutdown_reply"):
                if op == "connect_reply":
                    self._engines = list(msg["engines"]); self._engine_info = msg.get("engine_info", {})
                with self._reply_cv:
                    self._replies[op] = msg
                    self._reply_cv.notify_all()

    def _submit_many(self, kind, target_payloads, single, retries=0):
        ids = []
        now = time.time()
        for target, payload in target_payloads:
            tid = uuid.uuid4().hex
            self._tasks[tid] = {"done": threading.Event(), "status": "pending", "result": None, "error": None, "stdout": "", "stderr": "",
                                "data": {}, "submitted": now, "started": None, "completed": None, "engine_id": None if target == "lb" else target,
                                "requeues": 0}
            ids.append(tid)
            self._send({"op": "submit", "task_id": tid, "kind": kind, "target": target, "payload": payload, "retries": retries})
        return AsyncResult(self, ids, [t for t, _ in target_payloads], single)

    # ------------------------------------------------------------------ public API
    @property
    def ids(self):
        self._send({"op": "connect"})
        self._wait_reply("connect_reply", 5.0)
        return list(self._engines)

    def __len__(self):
        return len(self.ids)

    def wait_for_engines(self, n, timeout=60):
        """Block until ``n`` engines have registered (readiness handshake instead of ``sleep 30``)."""
        deadline = time.time() + timeout
        while time.time() < deadline:
            if len(self.ids) >= n:
                return self
            time.sleep(0.1)
        raise TimeoutError("only %d of %d engines registered after %s s" % (len(self._engines), n, timeout))

    def __getitem__(self, key):
        ids = self.ids
        if isinstance(key, slice):
            return DirectView(self, ids[key], False)
        if isinstance(key, (list, tuple)):
            return DirectView(self, list(key), False)
        if key not in ids:
            if isinstance(key, int) and -len(ids) <= key < 0:
                key = ids[key]
            else:
                raise IndexError("no engine with id %r" % (key,))
        return DirectView(self, [key], True)

    def direct_view(self, targets="all"):
        return self[:] if targets == "all" else self[targets]

    def load_balanced_view(self, targets=None):
        return LoadBalancedView(self, targets)

    def queue_status(self):
        self._send({"op": "status"})
        r = self._wait_reply("status_reply", 5.0)
        return {} if r is None else {"engines": r["engines"], "unassigned": r["lb_queued"]}

    def engine_info(self):
        _ = self.ids
        return dict(self._engine_info)

    def abort(self, jobs=None):
        for ar in (jobs if isinstance(jobs, (list, tuple)) else [jobs]):
            if ar is not None:
                ar.abort()

    def wait(self, jobs=None, timeout=None):
        ars = jobs if isinstance(jobs, (list, tuple)) else [jobs]
        return all(a.wait(timeout) for a in ars if a is not None)

    def shutdown(self, hub=True):
        """Stop the engines and the controller."""
        self._send({"op": "shutdown"})
        self._wait_reply("shutdown_reply", 3.0)
        self.close()

    def close(self):
        self._closed = True
        try:
            self._thread.join(1.0)
            self._sock.close(0)
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
