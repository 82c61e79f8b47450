#!/bin/bash
# Start a controller and one engine per GPU on this host (reference: salloc + ipcontroller + sleep 30 + srun ipengine).
#   ./startCluster.sh [n_engines] [slots_per_engine]
# Prints the cluster id to pass to Client(cluster_id=...).  Stop with: python -m cori_intml_examples_b200.cluster stop
. "$(dirname "$0")/setup.sh"
clusterID="b200_$$"
nEngines="${1:-}"
slots="${2:-1}"
echo "Launching controller and engines (cluster id $clusterID)"
exec python -m cori_intml_examples_b200.cluster start --cluster-id "$clusterID" --slots "$slots" ${nEngines:+-n "$nEngines"} --foreground
