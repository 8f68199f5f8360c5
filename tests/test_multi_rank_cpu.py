"""The N>1 path on CPU: sharding of camera-frames over ranks and the histogram sum across ranks (gloo, world 2)."""
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np

from multiview_b200 import shard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharding_covers_every_unit_exactly_once():
    for n_cams, n_frames in [(8, 1), (16, 300), (3, 10), (1, 7)]:
        for world in (1, 2, 4, 8):
            seen = {}
            for rank in range(world):
                for u in shard.units_for_rank(n_cams, n_frames, world, rank):
                    assert u not in seen, (u, rank, seen[u])
                    seen[u] = rank
            assert len(seen) == n_cams * n_frames, (n_cams, n_frames, world, len(seen))
    assert shard.cameras_for_rank(16, 8, 3) == [3, 11]  # camera-major: a camera never moves between GPUs


WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %r)
    import numpy as np, torch, torch.distributed as dist
    import oracle
    from multiview_b200 import shard, synth
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    n_cams, W, H = 4, 96, 64
    hp = oracle.HistParams(0.1, 0.5, 2.0, 0, 0)
    hw, hh = oracle.hist_dims(synth.HIST_AABB, 0.1)
    local = np.zeros((hh, hw), np.float32)
    hists = {}
    for cam in shard.cameras_for_rank(n_cams, world, rank):
        rng = np.random.default_rng(cam)
        disp = rng.uniform(2, 40, (H, W)).astype(np.float32)
        b = oracle.bcam_init(synth.bcam_quaternion(), synth.BCAM_T, synth.BCAM_BASELINE, synth.intrinsics(W, H))
        pc = oracle.point_cloud(b, disp)
        et = synth.camera_extrinsics(cam, n_cams)
        h = oracle.floor_hist_camera(hp, et, synth.HIST_AABB, et[:3], pc["Xs"], pc["xy"])
        hists[cam] = h
        local += h
    t = torch.from_numpy(local.copy())
    dist.all_reduce(t)                      # the only collective on the path (floor-histogram.cpp:296-300)
    np.save(os.path.join(%r, f"sum_{rank}.npy"), t.numpy())
    for cam, h in hists.items():
        np.save(os.path.join(%r, f"cam_{cam}.npy"), h)
    dist.destroy_process_group()
""")


def test_histogram_allreduce_equals_camera_order_sum(tmp_path):
    import oracle

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    script = tmp_path / "worker.py"
    script.write_text(WORKER % (ROOT, str(tmp_path), str(tmp_path)))
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env))
    for p in procs:
        assert p.wait(timeout=300) == 0
    cams = np.stack([np.load(tmp_path / f"cam_{c}.npy") for c in range(4)])
    want = oracle.hist_sum(cams)  # camera-order float sum; counts are integers < 2^24, so any order is exact
    for rank in range(2):
        assert np.array_equal(np.load(tmp_path / f"sum_{rank}.npy"), want)
    assert want.sum() > 100
