"""Multi-GPU host logic on CPU: spp-range sharding + one sum-reduce (gloo, world_size 2), with the oracle standing
in for the renderer.  The union of the shards must be the single-rank image (same global sample indices)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from rust_raytracer_b200.sharding import spp_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_spp_range_partitions_exactly():
    for world in (1, 2, 3, 4, 8):
        for total in (0, 1, 7, 8, 10000):
            cover = []
            for r in range(world):
                b, c = spp_range(r, world, total, spp_begin=5)
                cover += list(range(b, b + c))
            assert cover == list(range(5, 5 + total))
    with pytest.raises(ValueError):
        spp_range(2, 2, 10)


WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch, torch.distributed as dist
import rust_raytracer_b200 as rt
from rust_raytracer_b200.sharding import spp_range, reduce_accum
from oracle_ffi import OracleScene
from scenes_util import scene_builder
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
sb = scene_builder("cornell_book3")
orc = OracleScene(sb)
p = rt.copy_params(sb.params, width=24, height=24, spp_count=6)
begin, count = spp_range(rank, world, 6)
acc, _, _, _ = orc.render(sb.camera, rt.copy_params(p, spp_begin=begin, spp_count=count), n_jobs=4, n_threads=2)
t = torch.from_numpy(acc.copy())
reduce_accum(t, dst=0)
if rank == 0:
    full, _, _, _ = orc.render(sb.camera, p, n_jobs=4, n_threads=2)
    np.save(sys.argv[1], np.stack([t.numpy(), full]))
dist.destroy_process_group()
'''


def test_two_rank_reduce_equals_single_rank(tmp_path, built):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    out = tmp_path / "out.npy"
    env = dict(os.environ, OMP_NUM_THREADS="1")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29531",
                          str(script), str(out)], capture_output=True, text=True, timeout=300, env=env)
    assert res.returncode == 0, res.stderr[-2000:]
    both = np.load(out)
    np.testing.assert_allclose(both[0], both[1], rtol=1e-12, atol=1e-12)
    assert both[1].sum() > 0
