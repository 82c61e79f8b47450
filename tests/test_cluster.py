"""Control plane: controller + engines + client views + AsyncResult, thread and process engines."""
import functools
import time

import pytest

from cori_intml_examples_b200.cluster import LocalCluster, RemoteError, EngineError, TaskAborted


def _work(a, b=2, sleep=0.0):
    import os, time
    from cori_intml_examples_b200.cluster import publish_data
    print("pid", os.getpid())
    publish_data({"status": "running", "a": a})
    time.sleep(sleep)
    return a * b


@pytest.fixture(scope="module")
def proc_cluster():
    with LocalCluster(3, mode="process", gpus=[]) as cl:
        c = cl.client()
        yield cl, c
        c.close()


def test_ids_execute_get_push_pull(proc_cluster, capsys):
    cl, c = proc_cluster
    assert c.ids == [0, 1, 2] and len(c) == 3
    dv = c[:]
    dv.execute("import os\nx = 40 + engine_id\nprint('x is', x)")
    out = capsys.readouterr().out
    assert "[stdout:0] x is 40" in out and "[stdout:2] x is 42" in out       # %%px echo format
    assert dv.get("x") == [40, 41, 42] and c[1].get("x") == 41 and c[0]["x"] == 40
    dv.execute("class H: pass\nh = H(); h.epoch = [0, 1]; h.history = {'loss': [1.0, 0.5]}", silent=True)
    assert c[0].get("h.epoch") == [0, 1]                                      # attribute-path pull (DistTrain_rpv.ipynb:310)
    assert c[:].get("h.history")[2] == {"loss": [1.0, 0.5]}
    c[2]["y"] = {"k": 1}
    assert c[2].pull("y") == {"k": 1}
    dv.scatter("part", list(range(9)))
    assert dv.gather("part") == list(range(9))


def test_load_balanced_apply_and_async_result_surface(proc_cluster):
    cl, c = proc_cluster
    lv = c.load_balanced_view()
    ars = [lv.apply(_work, i, b=3, sleep=0.05) for i in range(7)] + [lv.apply(functools.partial(_work, b=5), 10)]
    assert [a.get(20) for a in ars] == [0, 3, 6, 9, 12, 15, 18, 50]
    ar = ars[0]
    assert ar.ready() and ar.done() and ar.successful()
    assert ar.stdout.startswith("pid ") and ar.stderr == ""
    assert ar.data == {"status": "running", "a": 0}
    assert ar.started <= ar.completed and ar.elapsed >= 0.04
    md = ar.metadata
    assert md["status"] == "ok" and md["engine_id"] in (0, 1, 2) and md["started"] is not None
    assert sum(a.ready() for a in ars) == len(ars)                            # DistHPO_mnist.ipynb:272 idiom
    assert {a.engine_id for a in ars} == {0, 1, 2}                            # least-load spread the tasks
    assert lv.map(lambda v: v + 1, [1, 2, 3]) == [2, 3, 4]


def test_interactive_functions_ship_by_value(proc_cluster):
    cl, c = proc_cluster
    offset = 7

    def closure(v):           # defined here, not importable on the engine: must travel by value
        import math
        return math.floor(v) + offset
    assert c.load_balanced_view().apply(closure, 2.5).get(10) == 9
    assert c[:].apply_sync(lambda: 1) == [1, 1, 1]


def test_remote_exception_and_published_data_while_running(proc_cluster):
    cl, c = proc_cluster
    lv = c.load_balanced_view()
    bad = lv.apply(lambda: 1 / 0)
    with pytest.raises(RemoteError) as ei:
        bad.get(10)
    assert ei.value.ename == "ZeroDivisionError" and "ZeroDivisionError" in ei.value.traceback
    assert not bad.successful()
    slow = lv.apply(_work, 5, sleep=0.6)
    t0 = time.time()
    while not slow.data and time.time() - t0 < 5:
        time.sleep(0.02)
    assert slow.data == {"status": "running", "a": 5} and not slow.ready()    # side channel before completion
    assert slow.get(10) == 10


def test_abort_running_and_queued(proc_cluster):
    cl, c = proc_cluster
    lv = c.load_balanced_view()

    def spin():
        import time
        for _ in range(2000):
            time.sleep(0.005)
        return "finished"
    running = [lv.apply(spin) for _ in range(3)]
    queued = lv.apply(spin)
    time.sleep(0.4)
    queued.abort(); running[0].abort()
    with pytest.raises(TaskAborted):
        queued.get(5)
    with pytest.raises(TaskAborted):
        running[0].get(5)
    for r in running[1:]:
        r.abort()
    c.wait(running, 5)
    assert c.queue_status()["unassigned"] == 0


def test_engine_death_fails_direct_and_requeues_load_balanced():
    import os
    os.environ["CIB_HEARTBEAT_TIMEOUT"] = "1.5"
    try:
        with LocalCluster(2, mode="process", gpus=[], env={"CIB_HEARTBEAT_TIMEOUT": "1.5", "CIB_HEARTBEAT_PERIOD": "0.3"}) as cl:
            c = cl.client()
            lv = c.load_balanced_view(); lv.retries = 1
            def slow(tag):
                import time
                time.sleep(1.0)
                return tag
            a = lv.apply(slow, "a"); b = lv.apply(slow, "b")
            time.sleep(0.3)
            victim = a.engine_id if a.engine_id is not None else 0
            cl.kill_engine(victim)                       # fault injection: SIGKILL, no goodbye
            assert sorted([a.get(15), b.get(15)]) == ["a", "b"]          # the lost task was re-queued on the survivor
            assert a.metadata["requeues"] + b.metadata["requeues"] == 1
            assert len(c.ids) == 1
            with pytest.raises(EngineError):
                c.load_balanced_view()                   # direct task to the dead engine fails instead of hanging
                from cori_intml_examples_b200.cluster.client import DirectView
                DirectView(c, [victim], True).apply_sync(lambda: 1)
            c.close()
    finally:
        os.environ.pop("CIB_HEARTBEAT_TIMEOUT", None)


def test_thread_engines_and_slots():
    with LocalCluster(1, mode="thread", slots=3) as cl:
        c = cl.client()
        lv = c.load_balanced_view()
        t0 = time.time()
        ars = [lv.apply(_work, i, sleep=0.4) for i in range(3)]
        assert [a.get(10) for a in ars] == [0, 2, 4]
        assert time.time() - t0 < 1.0                    # three trials ran concurrently on one engine (slots = 3)
        assert c.queue_status()["engines"][0]["slots"] == 3
        c.close()


def test_magics_functions(capsys):
    from cori_intml_examples_b200.cluster import magics
    c = magics.ipcluster("-n 2 --cpu -q interactive -t 30:00 -e CIB_TEST_FLAG=on,myenv -m json,collections")
    try:
        assert c.ids == [0, 1]
        magics.px("print('hello', engine_id)")
        out = capsys.readouterr().out
        assert "Slurm options -q -t are ignored" in out and "[stdout:1] hello 1" in out
        # -e reaches the engine processes' environment, -m modules are already imported in the engine namespace
        assert c[:].apply_sync(lambda: (__import__("os").environ.get("CIB_TEST_FLAG"), __import__("os").environ.get("CIB_ENV_NAME"))) == [("on", "myenv")] * 2
        assert c[:].get("json.__name__") == ["json", "json"]
    finally:
        magics.ipcluster("--stop")


def test_killed_engine_is_respawned_and_takes_new_work():
    """Elastic recovery: with respawn=True the launcher restarts an engine process that died; the replacement registers under
    a new engine id and the cluster is back to full strength."""
    with LocalCluster(2, mode="process", gpus=[], respawn=True,
                      env={"CIB_HEARTBEAT_TIMEOUT": "1.5", "CIB_HEARTBEAT_PERIOD": "0.3"}) as cl:
        c = cl.client()
        before = set(c.ids)
        assert len(before) == 2
        cl.kill_engine(sorted(before)[0])
        deadline = time.time() + 60
        while time.time() < deadline:
            ids = set(c.ids)
            if len(ids) == 2 and ids != before:
                break
            time.sleep(0.2)
        ids = set(c.ids)
        assert len(ids) == 2 and ids != before and cl.respawns == 1
        lv = c.load_balanced_view()
        res = [lv.apply(_work, i) for i in range(6)]
        assert sorted(r.get(30) for r in res) == [0, 2, 4, 6, 8, 10]
        assert {r.engine_id for r in res} <= ids
        c.close()
