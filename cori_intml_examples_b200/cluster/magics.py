"""``%ipcluster`` line magic and ``%%px`` cell magic (registered only when IPython is importable), plus the plain
functions they wrap so scripts and tests can use the same entry points without IPython.

``%ipcluster`` keeps the reference's docopt-style options (ipcluster_magics.py:14-34) -- ``-N/--num_nodes`` becomes
the number of GPUs/engines, ``-e`` / ``-m`` become engine environment variables / pre-imported modules, Slurm-only options (-C, -q, -J, -t, -d) are accepted and ignored with a note,
since there is no batch system on a single 8xB200 box.
"""
import shlex

from .launcher import LocalCluster
from .client import Client

_clusters = {}
_default_client = [None]

USAGE = """Usage: %ipcluster [-N <nodes>] [-n <engines>] [--slots <k>] [--cluster-id <id>] [--cpu] [--stop]
   -N, --num_nodes <int>   number of GPUs (one engine each)        [default: all]
   -n, --num_engines <int> number of engines                       [default: one per GPU]
   --slots <int>           concurrent trials per engine             [default: 1]
   --cluster-id <str>      name of the cluster                     [default: generated]
   --cpu                   CPU-only engines (no GPU pinning)
   --stop                  stop the cluster started from this session
   -e, --env <K=V,...>     environment variables for the engine processes (a bare name becomes CIB_ENV_NAME)
   -m, --modules <a,b>     Python modules imported in every engine at start-up
   -t -d -C -q -J          accepted for compatibility with the Cori version, ignored here
"""


def ipcluster(line=""):
    """Start (or stop) a local cluster; returns a connected Client."""
    args = shlex.split(line)
    opts = {"n": None, "slots": 1, "cluster_id": None, "cpu": False, "stop": False, "env": {}, "modules": []}
    ignored = []
    i = 0
    while i < len(args):
        a = args[i]
        if a in ("-N", "--num_nodes", "-n", "--num_engines"):
            opts["n"] = int(args[i + 1]); i += 2
        elif a == "--slots":
            opts["slots"] = int(args[i + 1]); i += 2
        elif a == "--cluster-id":
            opts["cluster_id"] = args[i + 1]; i += 2
        elif a == "--cpu":
            opts["cpu"] = True; i += 1
        elif a == "--stop":
            opts["stop"] = True; i += 1
        elif a in ("-e", "--env"):
            # the reference activates a conda environment of that name in the engines' job script (ipcluster_magics.py:22); here the
            # engines are local processes: KEY=VALUE[,KEY=VALUE...] goes into their environment, a bare name is kept as CIB_ENV_NAME
            for item in args[i + 1].split(","):
                if "=" in item:
                    k, v = item.split("=", 1); opts["env"][k] = v
                elif item:
                    opts["env"]["CIB_ENV_NAME"] = item
            i += 2
        elif a in ("-m", "--modules"):
            # modules to load on Cori = Python modules pre-imported in every engine's namespace here (comma separated)
            opts["modules"] += [m for m in args[i + 1].split(",") if m]; i += 2
        elif a in ("-t", "--time", "-d", "--dir", "-C", "--const", "-q", "--queue", "-J", "--name"):
            ignored.append(a); i += 2
        elif a in ("-h", "--help"):
            print(USAGE); return None
        else:
            raise ValueError("unknown %%ipcluster option %r\n%s" % (a, USAGE))
    if opts["stop"]:
        for c in list(_clusters.values()):
            c.stop()
        _clusters.clear(); _default_client[0] = None
        return None
    if ignored:
        print("note: Slurm options %s are ignored (single-host cluster, no batch system)" % " ".join(ignored))
    cl = LocalCluster(opts["n"], opts["cluster_id"], opts["slots"], gpus=[] if opts["cpu"] else None, env=opts["env"]).start()
    _clusters[cl.cluster_id] = cl
    client = cl.client()
    for mod in opts["modules"]:
        try:
            client[:].execute("import %s" % mod, block=True, silent=True)
        except Exception as e:                     # a module that does not exist on this box must not kill the cluster start
            print("note: engines could not import %r: %s" % (mod, e))
    _default_client[0] = client
    print("cluster_id = %r, engines = %s" % (cl.cluster_id, client.ids))
    return client


def px(code, client=None, targets=None):
    """Run ``code`` on all (or the given) engines and echo their output as ``[stdout:i] ...`` (the ``%%px`` magic)."""
    client = client or _default_client[0]
    if client is None:
        raise RuntimeError("no client: call ipcluster(...) or pass client=Client(...)")
    view = client[:] if targets is None else client[targets]
    return view.execute(code, block=True)


class IPClusterMagics(object):
    """Class form of the magics, same name as the reference's (ipcluster_magics.py:12): ``IPClusterMagics().ipcluster(line)``
    starts a cluster and returns the client; its docstring is the usage text, like the reference's docopt class docstring."""
    __doc__ = (__doc__ or "") + "\n\n" + USAGE

    def __init__(self, shell=None):
        self.shell = shell

    def ipcluster(self, line=""):
        return ipcluster(line)

    def px(self, line="", cell=""):
        return px(cell or line)


def set_default_client(client):
    _default_client[0] = client


def load_ipython_extension(ip):      # pragma: no cover -- needs IPython
    from IPython.core.magic import register_line_magic, register_cell_magic

    @register_line_magic("ipcluster")
    def _ipcluster(line):
        return ipcluster(line)

    @register_cell_magic("px")
    def _px(line, cell):
        px(cell)


try:                                   # pragma: no cover
    from IPython import get_ipython
    if get_ipython() is not None:
        load_ipython_extension(get_ipython())
except Exception:
    pass
