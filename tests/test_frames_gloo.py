"""CPU, world_size 2 over gloo: frame sharding and the summary all_gather used when trajectory
frames are partitioned across GPUs.  The data path has no collective;
ranks exchange only fixed-size per-frame summary records."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, nframes, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from memsurfer_b200.frame import shard_frames, gather_summaries
    mine = shard_frames(nframes, rank, world)
    rows = []
    nsteps = (nframes + world - 1) // world
    for step in range(nsteps):
        # a rank that ran out of frames contributes a NaN record so that every rank calls the collective
        f = mine[step] if step < len(mine) else -1
        rec = torch.full((2, 3, 4), float("nan"), dtype=torch.float64)
        if f >= 0:
            rec = torch.arange(24, dtype=torch.float64).reshape(2, 3, 4) + 1000.0 * f     # stand-in for (leaflet, field, 4)
        g = gather_summaries(rec)                                                          # (W, 2, 3, 4)
        assert g.shape == (world, 2, 3, 4)
        for r in range(world):
            fr = step * world + r
            if fr < nframes:
                assert torch.equal(g[r], torch.arange(24, dtype=torch.float64).reshape(2, 3, 4) + 1000.0 * fr)
                rows.append(fr)
            else:
                assert torch.isnan(g[r]).all()
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), np.array(rows))
    dist.destroy_process_group()


@pytest.mark.parametrize("nframes", [4, 5])
def test_frame_sharding_and_summary_gather_world2(tmp_path, nframes):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), nframes, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        seen = np.load(tmp_path / f"rank{r}.npy")
        assert sorted(seen.tolist()) == list(range(nframes))          # every rank ends with every frame's record


def test_shard_frames_partitions_exactly():
    from memsurfer_b200.frame import shard_frames
    for n in (0, 1, 7, 1000):
        for w in (1, 2, 4, 8):
            parts = [shard_frames(n, r, w) for r in range(w)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_gather_without_process_group_is_identity():
    from memsurfer_b200.frame import gather_summaries
    x = torch.arange(6, dtype=torch.float64).reshape(2, 3)
    assert torch.equal(gather_summaries(x), x.unsqueeze(0))
