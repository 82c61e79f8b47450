#!/usr/bin/env python
"""torchrun probe: can this box bind per-process gradient buffers into one NVSwitch multicast object?  Prints the outcome per rank."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
os.environ["CIB_NVLS"] = "1"; os.environ["CIB_NVLS_DEBUG"] = "1"
import torch
from cori_intml_examples_b200.models import rpv
from cori_intml_examples_b200.parallel import dist as hvd
hvd.init()
dev = hvd.local_rank() % torch.cuda.device_count(); torch.cuda.set_device(dev)
m = rpv.build_model((64, 64, 3), [16, 32, 64], [128], dropout=0.2, optimizer="Adam", lr=1e-3, use_horovod=True, device="cuda:%d" % dev)
m.backend._ensure(32)
print("rank %d: nvls active = %s" % (hvd.rank(), bool(m.backend.lib.cib_engine_nvls_active(m.backend.engine))), flush=True)
hvd.shutdown()
