"""CPU (gloo, world_size 2 and 3): the slab / ghost / edge / replicated-union-find / global-order protocol of the multi-GPU
path (csrc/multigpu.cu), modelled in tests/dist/dist_model.py with the oracle as the local linker, reproduces the
single-domain catalogue exactly."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world,port", [(2, 29541), (3, 29542)])
def test_protocol_reproduces_single_domain_catalogue(world, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "dist", "dist_model.py"), "20"]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    print(p.stdout[-3000:])
    assert p.returncode == 0
    assert "-> OK" in p.stdout
