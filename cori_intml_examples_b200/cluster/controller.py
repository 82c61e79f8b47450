"""Controller: registers engines, routes direct tasks, load-balances queued tasks, relays results / published data /
output streams to the submitting client, watches heartbeats and requeues the tasks of dead engines.

Replaces ``ipcontroller`` (reference startCluster.sh:14; ipcluster_magics.py:75-79).  Scheduling matches the
IPyParallel default the reference relies on -- least load with a high-water-mark of one outstanding task per engine
 -- but the mark is per *slot*: an engine that advertises ``slots = k`` (k concurrent CUDA-graph
trials on one GPU) receives up to k tasks.

Run:  python -m cori_intml_examples_b200.cluster.controller --cluster-id ID [--ip 127.0.0.1]
"""
import argparse
import collections
import os
import secrets
import signal
import sys
import time

import zmq

from . import protocol as P

HEARTBEAT_TIMEOUT = float(os.environ.get("CIB_HEARTBEAT_TIMEOUT", "6.0"))


class EngineInfo(object):
    def __init__(self, ident, eid, info):
        self.ident, self.id, self.info = ident, eid, info
        self.slots = int(info.get("slots", 1))
        self.running = {}            # task_id -> task
        self.queue = collections.deque()   # direct tasks waiting for a slot (execute/push/pull keep order)
        self.last_seen = time.time()
        self.alive = True


class Controller(object):
    def __init__(self, cluster_id, ip="127.0.0.1", log=None):
        self.cluster_id = cluster_id
        self.key = secrets.token_bytes(32)
        self.ctx = zmq.Context.instance()
        self.client_sock = self.ctx.socket(zmq.ROUTER)
        self.engine_sock = self.ctx.socket(zmq.ROUTER)
        cport = self.client_sock.bind_to_random_port("tcp://%s" % ip)
        eport = self.engine_sock.bind_to_random_port("tcp://%s" % ip)
        self.info = {"cluster_id": cluster_id, "client_url": "tcp://%s:%d" % (ip, cport), "engine_url": "tcp://%s:%d" % (ip, eport),
                     "key": self.key.hex(), "pid": os.getpid(), "started": time.time()}
        self.engines = {}            # id -> EngineInfo
        self.by_ident = {}
        self.next_engine_id = 0
        self.lb_queue = collections.deque()
        self.tasks = {}              # task_id -> task dict (client ident, payload, ...)
        self.running = True
        self.log = log or (lambda *a: None)

    # ------------------------------------------------------------------ helpers
    def send_client(self, ident, msg):
        self.client_sock.send_multipart([ident] + P.pack(msg, self.key))

    def send_engine(self, eng, msg):
        self.engine_sock.send_multipart([eng.ident] + P.pack(msg, self.key))

    def alive_ids(self):
        return sorted(e.id for e in self.engines.values() if e.alive)

    def dispatch(self, eng, task):
        task["engine_id"] = eng.id
        task["state"] = "running"
        eng.running[task["task_id"]] = task
        self.send_engine(eng, {"op": "task", "task_id": task["task_id"], "kind": task["kind"], "payload": task["payload"]})

    def schedule(self):
        # direct queues first (per-engine order), then the load-balanced queue: least-loaded engine with a free slot
        for eng in self.engines.values():
            while eng.alive and eng.queue and len(eng.running) < eng.slots:
                self.dispatch(eng, eng.queue.popleft())
        while self.lb_queue:
            free = [e for e in self.engines.values() if e.alive and len(e.running) + len(e.queue) < e.slots]
            if not free:
                break
            eng = min(free, key=lambda e: (len(e.running) / float(e.slots), e.id))
            self.dispatch(eng, self.lb_queue.popleft())

    def finish(self, task, msg):
        task["state"] = "done"
        self.tasks.pop(task["task_id"], None)
        self.send_client(task["client"], msg)

    # ------------------------------------------------------------------ client messages
    def on_client(self, ident, msg):
        op = msg["op"]
        if op == "connect":
            self.send_client(ident, {"op": "connect_reply", "engines": self.alive_ids(), "cluster_id": self.cluster_id,
                                     "engine_info": {e.id: e.info for e in self.engines.values() if e.alive}})
        elif op == "submit":
            task = {"task_id": msg["task_id"], "kind": msg["kind"], "payload": msg["payload"], "client": ident,
                    "retries": int(msg.get("retries", 0)), "target": msg["target"], "state": "queued", "engine_id": None,
                    "submitted": time.time()}
            self.tasks[task["task_id"]] = task
            if msg["target"] == "lb":
                if not self.alive_ids():
                    self.finish(task, self._error_msg(task, "EngineError", "no engines registered", None))
                    return
                self.lb_queue.append(task)
            else:
                eng = self.engines.get(msg["target"])
                if eng is None or not eng.alive:
                    self.finish(task, self._error_msg(task, "EngineError", "engine %r is not registered" % (msg["target"],), msg["target"]))
                    return
                eng.queue.append(task)
            self.schedule()
        elif op == "abort":
            task = self.tasks.get(msg["task_id"])
            if task is None:
                return
            if task["state"] == "queued":
                for q in [self.lb_queue] + [e.queue for e in self.engines.values()]:
                    try:
                        q.remove(task)
                    except ValueError:
                        pass
                self.finish(task, self._error_msg(task, "TaskAborted", "aborted before start", None, aborted=True))
            else:
                task["aborting"] = True
                eng = self.engines.get(task["engine_id"])
                if eng is not None and eng.alive:
                    self.send_engine(eng, {"op": "abort", "task_id": task["task_id"]})
        elif op == "status":
            self.send_client(ident, {"op": "status_reply", "req": msg.get("req"),
                                     "engines": {e.id: {"running": list(e.running), "queued": len(e.queue), "slots": e.slots, "alive": e.alive,
                                                        "info": e.info} for e in self.engines.values()},
                                     "lb_queued": len(self.lb_queue)})
        elif op == "shutdown":
            for eng in self.engines.values():
                if eng.alive:
                    self.send_engine(eng, {"op": "shutdown"})
            self.send_client(ident, {"op": "shutdown_reply"})
            self.running = False

    def _error_msg(self, task, ename, evalue, engine_id, aborted=False):
        now = time.time()
        return {"op": "result", "task_id": task["task_id"], "status": "aborted" if aborted else "error", "engine_id": engine_id,
                "error": (ename, evalue, ""), "stdout": "", "stderr": "", "started": task.get("started"), "completed": now, "payload": None}

    # ------------------------------------------------------------------ engine messages
    def on_engine(self, ident, msg):
        op = msg["op"]
        eng = self.by_ident.get(ident)
        if op == "register":
            eid = self.next_engine_id; self.next_engine_id += 1
            eng = EngineInfo(ident, eid, msg.get("info", {}))
            self.engines[eid] = eng; self.by_ident[ident] = eng
            self.send_engine(eng, {"op": "registered", "engine_id": eid, "cluster_id": self.cluster_id})
            self.log("engine %d registered: %s" % (eid, eng.info))
            self.schedule()
            return
        if eng is None:
            return
        eng.last_seen = time.time()
        if op == "heartbeat":
            return
        task = self.tasks.get(msg.get("task_id"))
        if op == "started" and task is not None:
            task["started"] = msg["time"]
            self.send_client(task["client"], dict(msg, engine_id=eng.id))
        elif op in ("data", "stream") and task is not None:
            self.send_client(task["client"], dict(msg, engine_id=eng.id))
        elif op == "result":
            eng.running.pop(msg["task_id"], None)
            if task is not None:
                self.finish(task, dict(msg, engine_id=eng.id))
            self.schedule()
        elif op == "unregister":
            self.engine_lost(eng, "engine unregistered")

    def engine_lost(self, eng, why):
        if not eng.alive:
            return
        eng.alive = False
        self.log("engine %d lost: %s" % (eng.id, why))
        for task in list(eng.running.values()) + list(eng.queue):
            if task["target"] == "lb" and task["retries"] > 0 and not task.get("aborting"):
                task["retries"] -= 1; task["state"] = "queued"; task["engine_id"] = None
                self.lb_queue.appendleft(task)
                self.send_client(task["client"], {"op": "requeued", "task_id": task["task_id"], "engine_id": eng.id})
            else:
                self.finish(task, self._error_msg(task, "EngineError", "engine %d died: %s" % (eng.id, why), eng.id))
        eng.running.clear(); eng.queue.clear()
        self.schedule()

    # ------------------------------------------------------------------ main loop
    def serve(self):
        P.write_connection_file(self.cluster_id, self.info)
        poller = zmq.Poller()
        poller.register(self.client_sock, zmq.POLLIN)
        poller.register(self.engine_sock, zmq.POLLIN)
        try:
            while self.running:
                for sock, _ in poller.poll(250):
                    frames = sock.recv_multipart()
                    try:
                        msg = P.unpack(frames, self.key)
                    except Exception:
                        continue
                    if sock is self.client_sock:
                        self.on_client(frames[0], msg)
                    else:
                        self.on_engine(frames[0], msg)
                now = time.time()
                for eng in self.engines.values():
                    if eng.alive and now - eng.last_seen > HEARTBEAT_TIMEOUT:
                        self.engine_lost(eng, "no heartbeat for %.1f s" % (now - eng.last_seen))
        finally:
            time.sleep(0.1)
            try:
                os.remove(P.connection_file(self.cluster_id))
            except OSError:
                pass
            self.client_sock.close(0); self.engine_sock.close(0)


def main(argv=None):
    ap = argparse.ArgumentParser(description="cib200 cluster controller")
    ap.add_argument("--cluster-id", default="cib_%d" % os.getpid())
    ap.add_argument("--ip", default="127.0.0.1")
    ap.add_argument("--log-to-file", action="store_true")
    a = ap.parse_args(argv)
    logf = open(os.path.join(P.cluster_dir(), "controller-%s.log" % a.cluster_id), "a") if a.log_to_file else sys.stderr

    def log(*args):
        logf.write("[controller %s] %s\n" % (time.strftime("%H:%M:%S"), " ".join(str(x) for x in args))); logf.flush()
    c = Controller(a.cluster_id, a.ip, log)
    signal.signal(signal.SIGTERM, lambda *_: setattr(c, "running", False))
    log("listening: clients %s, engines %s" % (c.info["client_url"], c.info["engine_url"]))
    c.serve()


if __name__ == "__main__":
    main()
