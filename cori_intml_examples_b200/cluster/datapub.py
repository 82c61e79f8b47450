"""Engine -> client side channel: ``publish_data(dict)`` replaces the running task's ``AsyncResult.data`` on the client.

Same role as ``ipyparallel.datapub.publish_data`` (reference mlextras.py:6,21-33).  Outside an engine task (plain
script, unit test) the data goes to a process-local sink so callbacks using it keep working.
"""
import threading

_local = threading.local()
_sink = []


def _set_current(ctx):
    _local.ctx = ctx


def current_task():
    return getattr(_local, "ctx", None)


def publish_data(data):
    ctx = current_task()
    if ctx is None:
        _sink.append(data)
        if len(_sink) > 64:
            del _sink[:-64]
        return
    ctx.publish(data)


def last_published():
    """What a non-engine caller published last (testing aid)."""
    return _sink[-1] if _sink else None


def abort_requested():
    """True once the client asked to abort the running task (polled by training callbacks between epochs)."""
    ctx = current_task()
    return bool(ctx is not None and ctx.abort.is_set())
