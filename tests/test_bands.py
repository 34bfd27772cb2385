"""Row-band sharding on the CPU: the host statement of the library's RowMap, and the N > 1 gather run
with world_size 2 over gloo - each rank renders ITS bands of a golden scene with the oracle (the oracle is the
checker here; on the GPU box the same logic runs with libgmlrt and NCCL / IPC stores, see bench.py) and rank 0 must
end up with the reference's golden image byte for byte."""
import os
import socket
import sys

import numpy as np
import pytest

from go_raytracer_b200 import bands

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("height", [1, 7, 8, 9, 77, 150, 2160, 4320])
@pytest.mark.parametrize("n_parts", [1, 2, 3, 4, 8])
def test_parts_partition_the_rows(height, n_parts):
    seen = np.zeros(height, int)
    for part in range(n_parts):
        rows = bands.rows_of(part, n_parts, height)
        assert len(rows) == bands.n_rows(part, n_parts, height)
        assert (rows < height).all() and (np.diff(rows) > 0).all()              # in range, rendered top to bottom
        assert all(bands.owner_of(int(y), n_parts) == part for y in rows)
        seen[rows] += 1
    assert (seen == 1).all()
    if height >= 8 * bands.BAND_ROWS * n_parts:                                 # balance: within one band of each other
        sizes = [bands.n_rows(p, n_parts, height) for p in range(n_parts)]
        assert max(sizes) - min(sizes) <= bands.BAND_ROWS


@pytest.mark.parametrize("height", [1, 7, 8, 9, 77, 150, 2160])
@pytest.mark.parametrize("n_parts", [1, 2, 3, 8])
def test_row_blocks_of_the_host_gather_are_exactly_the_owned_rows(height, n_parts):
    """gmlrt_download_rows copies a part's rows as one strided 2-D copy + a tail: together the parts rebuild the frame."""
    src = np.arange(height * 3, dtype=np.int64).reshape(height, 3) + 1
    dst = np.zeros_like(src)
    for part in range(n_parts):
        one = bands.copy_own_rows(np.zeros_like(src), src, part, n_parts)
        rows = bands.rows_of(part, n_parts, height)
        assert (one[rows] == src[rows]).all() and one.sum() == src[rows].sum()      # its rows and nothing else
        bands.copy_own_rows(dst, src, part, n_parts)
    assert np.array_equal(dst, src)


def test_row_map_matches_the_header_constants():
    src = open(os.path.join(HERE, "..", "go-raytracer_b200", "csrc", "gmlrt_api.cu")).read()
    assert "c->band_rows = nd > 1 ? %d : c->scene.height" % bands.BAND_ROWS in src


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.join(HERE, ".."))
    import torch
    import torch.distributed as dist
    import oracle
    from util import scene_blobs
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        blob = scene_blobs("sphere")[0][1]
        import struct
        w, h = struct.unpack_from("<ii", blob, 16)
        frame = np.zeros((h, w, 4), np.uint8)
        rows = bands.rows_of(rank, world, h)
        # contiguous runs of owned rows = bands; the oracle renders a crop with the reference's per-(column, 20-row batch) PCG streams
        for b0 in range(0, len(rows), bands.BAND_ROWS):
            y0, y1 = int(rows[b0]), int(rows[min(b0 + bands.BAND_ROWS, len(rows)) - 1]) + 1
            img, _ = oracle.render(blob, threads=2, crop=(0, y0, w, y1))
            frame[y0:y1] = img[y0:y1] if img.shape[0] == h else img
        t = torch.from_numpy(frame)
        bands.gather_sum(t, dist, dst=0)
        counts = torch.tensor([len(rows)])
        dist.all_reduce(counts)
        if rank == 0:
            assert int(counts[0]) == h
            np.save(out_path, t.numpy())
    finally:
        dist.destroy_process_group()


def _worker_shared(rank, world, port, path, h, w):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, os.path.join(HERE, ".."))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shared = np.memmap(path, dtype=np.uint8, mode="r+", shape=(h, w, 4))
        mine = np.zeros((h, w, 4), np.uint8)                     # this rank's "device frame": only its own rows are valid
        rows = bands.rows_of(rank, world, h)
        mine[rows] = (rows[:, None, None] * 7 + np.arange(w)[None, :, None] + np.arange(4)[None, None, :]) % 251
        bands.copy_own_rows(shared, mine, rank, world)           # what gmlrt_download_rows does over this rank's PCIe link
        shared.flush()
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_ranks_fill_one_shared_host_raster(tmp_path):
    """The host gather of bench.py (N > 1): every rank writes only its bands into one raster all ranks map; after the
    barrier the raster is complete - no collective on the data path."""
    import torch.multiprocessing as mp
    h, w = 77, 12
    path = str(tmp_path / "frame.bin")
    with open(path, "wb") as f:
        f.truncate(h * w * 4)
    mp.spawn(_worker_shared, args=(2, _free_port(), path, h, w), nprocs=2, join=True)
    got = np.memmap(path, dtype=np.uint8, mode="r", shape=(h, w, 4))
    y = np.arange(h)
    want = ((y[:, None, None] * 7 + np.arange(w)[None, :, None] + np.arange(4)[None, None, :]) % 251).astype(np.uint8)
    assert np.array_equal(np.asarray(got), want)


def test_two_ranks_gather_the_golden_over_gloo(tmp_path):
    import torch.multiprocessing as mp
    from util import golden
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    assert np.array_equal(got, golden("example_sphere.png"))
