"""Engine: a long-lived worker process (one per GPU) with a persistent namespace, executing tasks sent by the controller.

Replaces ``ipengine`` (reference startCluster.sh:18; ipcluster_magics.py:70-73).  Task kinds: ``execute`` (run source
in the persistent namespace = ``%%px``), ``apply`` (call a by-value pickled function), ``pull`` (evaluate an
expression such as ``history.epoch``, DistTrain_rpv.ipynb:310), ``push``.  Per-task stdout/stderr are captured and
returned; ``publish_data`` (see datapub.py) streams dicts to the client while the task runs; ``abort`` raises
``TaskAborted`` inside the running task.  ``--slots k`` runs up to k ``apply`` tasks concurrently in threads that
share this process' CUDA context, so k small CUDA-graph-captured trials overlap on one GPU.

Run:  python -m cori_intml_examples_b200.cluster.engine --cluster-id ID [--gpu N] [--slots K]
"""
import argparse
import ctypes
import io
import os
import queue
import socket
import sys
import threading
import time
import traceback

import zmq

from . import protocol as P
from . import datapub

HEARTBEAT_PERIOD = float(os.environ.get("CIB_HEARTBEAT_PERIOD", "1.0"))


class _ThreadStream(io.TextIOBase):
    """sys.stdout / sys.stderr replacement that routes writes to the running task of the calling thread."""

    def __init__(self, name, fallback):
        self._name, self._fallback = name, fallback

    def write(self, s):
        ctx = datapub.current_task()
        if ctx is None:
            return self._fallback.write(s)
        ctx.write(self._name, s)
        return len(s)

    def flush(self):
        ctx = datapub.current_task()
        if ctx is None:
            self._fallback.flush()

    def isatty(self):
        return False


class TaskContext(object):
    def __init__(self, engine, task_id):
        self.engine, self.task_id = engine, task_id
        self.out = {"stdout": io.StringIO(), "stderr": io.StringIO()}
        self.abort = threading.Event()
        self.thread_id = None
        self.guard = threading.Lock()          # serialises abort injection against the task leaving its try block
        self.in_user_code = False
        self._last_stream = 0.0
        self._sent = {"stdout": 0, "stderr": 0}

    def write(self, name, s):
        self.out[name].write(s)
        now = time.time()
        if now - self._last_stream > 0.5:          # incremental output for AsyncResult.stdout while the task runs
            self._last_stream = now
            self.flush_streams()

    def flush_streams(self):
        for name in ("stdout", "stderr"):
            text = self.out[name].getvalue()
            if len(text) > self._sent[name]:
                self.engine.outbox.put({"op": "stream", "task_id": self.task_id, "name": name, "text": text[self._sent[name]:]})
                self._sent[name] = len(text)

    def publish(self, data):
        self.engine.outbox.put({"op": "data", "task_id": self.task_id, "data": P.dumps(data)})


class Engine(object):
    def __init__(self, cluster_id, gpu=None, slots=1, timeout=60.0, in_process=False):
        self.cluster_id, self.gpu, self.slots, self.in_process = cluster_id, gpu, int(slots), in_process
        info = P.read_connection_file(cluster_id, timeout)
        self.key = bytes.fromhex(info["key"])
        self.ctx = zmq.Context.instance()
        self.sock = self.ctx.socket(zmq.DEALER)
        self.sock.connect(info["engine_url"])
        self.outbox = queue.Queue()
        self.inbox = queue.Queue()
        self.ns = {"__name__": "__engine__", "__builtins__": __builtins__}
        self.running = True
        self.id = None
        self.contexts = {}
        self.exec_lock = threading.Lock()      # execute/push/pull mutate the shared namespace: one at a time

    def send(self, msg):
        self.sock.send_multipart(P.pack(msg, self.key))

    # ------------------------------------------------------------------ task execution (worker threads)
    def worker(self):
        while self.running:
            try:
                msg = self.inbox.get(timeout=0.2)
                if msg is None:
                    break
                self.run_task(msg)
            except queue.Empty:
                continue
            except P.TaskAborted:               # a late asynchronous abort must never kill the slot's worker thread
                continue

    def run_task(self, msg):
        tid, kind = msg["task_id"], msg["kind"]
        ctx = TaskContext(self, tid)
        ctx.thread_id = threading.get_ident()
        self.contexts[tid] = ctx
        datapub._set_current(ctx)
        started = time.time()
        self.outbox.put({"op": "started", "task_id": tid, "time": started})
        status, payload, error = "ok", None, None
        try:
            if ctx.abort.is_set():
                raise P.TaskAborted("TaskAborted", "aborted", "")
            with ctx.guard:
                ctx.in_user_code = True          # from here to the `finally` below an injected TaskAborted lands inside this try block
            body = P.loads(msg["payload"])
            if kind == "apply":
                f, args, kwargs = body
                payload = f(*args, **kwargs)
            elif kind == "execute":
                with self.exec_lock:
                    exec(compile(body, "<px>", "exec"), self.ns)
            elif kind == "pull":
                with self.exec_lock:
                    payload = [eval(e, self.ns) for e in body] if isinstance(body, (list, tuple)) else eval(body, self.ns)
            elif kind == "push":
                with self.exec_lock:
                    self.ns.update(body)
            else:
                raise ValueError("unknown task kind %r" % kind)
            payload = P.dumps(payload)
        except P.TaskAborted:
            status, error = "aborted", ("TaskAborted", "task aborted on engine %s" % self.id, "")
        except BaseException as e:      # noqa: B902 -- report everything, keep the engine alive
            status = "error"
            error = (type(e).__name__, str(e), traceback.format_exc())
            if isinstance(e, (KeyboardInterrupt, SystemExit)) and not self.in_process:
                self.running = False
        finally:
            # close the injection window BEFORE leaving the try block's protection: abort_task only injects while in_user_code is set
            # and holds the same lock, so an asynchronous TaskAborted can no longer land in the code below (result hand-off, the
            # worker loop) where nothing would catch it
            try:
                with ctx.guard:
                    ctx.in_user_code = False
            except P.TaskAborted:
                with ctx.guard:
                    ctx.in_user_code = False
                if status == "ok":
                    status, error = "aborted", ("TaskAborted", "task aborted on engine %s" % self.id, "")
            datapub._set_current(None)
            self.contexts.pop(tid, None)
        self.outbox.put({"op": "result", "task_id": tid, "status": status, "payload": payload if status == "ok" else None, "error": error,
                         "stdout": ctx.out["stdout"].getvalue(), "stderr": ctx.out["stderr"].getvalue(),
                         "started": started, "completed": time.time()})

    def abort_task(self, tid):
        ctx = self.contexts.get(tid)
        if ctx is None:
            return
        ctx.abort.set()
        with ctx.guard:
            if ctx.thread_id is not None and ctx.in_user_code:
                # asynchronous exception: lands at the next Python bytecode of the task's thread, which is inside run_task's try block
                # for as long as in_user_code is set (a task that has already left it just finishes normally)
                ctypes.pythonapi.PyThreadState_SetAsyncExc(ctypes.c_ulong(ctx.thread_id), ctypes.py_object(P.TaskAborted))

    # ------------------------------------------------------------------ I/O loop (owns the socket)
    def serve(self):
        info = {"pid": os.getpid(), "host": socket.gethostname(), "gpu": self.gpu, "slots": self.slots,
                "cuda_visible_devices": os.environ.get("CUDA_VISIBLE_DEVICES")}
        self.send({"op": "register", "info": info})
        if not self.in_process:
            sys.stdout = _ThreadStream("stdout", sys.__stdout__)
            sys.stderr = _ThreadStream("stderr", sys.__stderr__)
        workers = [threading.Thread(target=self.worker, daemon=True, name="cib-engine-worker-%d" % i) for i in range(self.slots)]
        for w in workers:
            w.start()
        poller = zmq.Poller(); poller.register(self.sock, zmq.POLLIN)
        last_hb = 0.0
        while self.running:
            for _ in poller.poll(20):
                try:
                    msg = P.unpack(self.sock.recv_multipart(), self.key)
                except Exception:
                    continue
                op = msg["op"]
                if op == "registered":
                    self.id = msg["engine_id"]
                    self.ns["engine_id"] = self.id
                elif op == "task":
                    self.inbox.put(msg)
                elif op == "abort":
                    self.abort_task(msg["task_id"])
                elif op == "shutdown":
                    self.running = False
            while True:
                try:
                    self.send(self.outbox.get_nowait())
                except queue.Empty:
                    break
            now = time.time()
            if now - last_hb > HEARTBEAT_PERIOD:
                self.send({"op": "heartbeat"}); last_hb = now
        for _ in workers:
            self.inbox.put(None)
        try:
            self.send({"op": "unregister"})
        except Exception:
            pass
        time.sleep(0.05)
        self.sock.close(0)


def main(argv=None):
    ap = argparse.ArgumentParser(description="cib200 cluster engine")
    ap.add_argument("--cluster-id", required=True)
    ap.add_argument("--gpu", type=int, default=None, help="GPU ordinal this engine owns (sets CUDA_VISIBLE_DEVICES)")
    ap.add_argument("--shared-visibility", action="store_true",
                    help="keep every GPU visible and only select --gpu (needed for data-parallel jobs spanning engines: "
                         "their peer-to-peer kernels map the other engines' memory)")
    ap.add_argument("--slots", type=int, default=1, help="concurrent apply-tasks (trials) on this engine's GPU")
    ap.add_argument("--timeout", type=float, default=60.0)
    a = ap.parse_args(argv)
    if a.gpu is not None:
        if a.shared_visibility:
            os.environ["CIB_DEVICE"] = str(a.gpu)
        elif "CUDA_VISIBLE_DEVICES" not in os.environ:
            os.environ["CUDA_VISIBLE_DEVICES"] = str(a.gpu)
    Engine(a.cluster_id, a.gpu, a.slots, a.timeout).serve()


if __name__ == "__main__":
    main()
