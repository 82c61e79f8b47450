"""Cluster control plane: controller + engines (one per GPU) + client views, over ZeroMQ on one host.

Replaces the IPyParallel layer of the reference (startCluster.sh, ipcluster_magics.py, ``ipp.Client`` usage in
every Dist*.ipynb) -- see SURVEY 2.4.3 for the surface.
"""
from .client import Client, AsyncResult, DirectView, LoadBalancedView
from .datapub import publish_data
from .launcher import LocalCluster, start_cluster, stop_cluster
from .protocol import RemoteError, EngineError, TaskAborted

__all__ = ["Client", "AsyncResult", "DirectView", "LoadBalancedView", "publish_data", "LocalCluster", "start_cluster",
           "stop_cluster", "RemoteError", "EngineError", "TaskAborted"]
