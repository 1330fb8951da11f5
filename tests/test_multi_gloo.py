"""N > 1 host logic on CPU (gloo, world_size 2): every rank renders the pass samples s with s mod N == rank
(here through the CPU emulation of the stages), the pass films are combined with ONE all-reduce(sum), and the
result equals the single-rank film -- the same flow bench.py / a multi-GPU host runs with NCCL."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers
from libyafaray_b200 import interface, scenes_pt

KW = dict(width=40, height=24, spp=5, bounces=3)


def _assemble():
    yi = interface.Interface()
    scenes_pt.cornell_box(yi, accelerator="yafaray-b200-bvh8", **KW)
    return yi.assemble()


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w, rgba = helpers.emu_pt_render(_assemble(), shard_index=rank, shard_count=world)
    film = torch.from_numpy(np.concatenate([w.reshape(-1), rgba.reshape(-1)]))
    dist.all_reduce(film)                       # one sum per pass (ImageFilm merge semantics, imagefilm.cc:940-1018)
    if rank == 0:
        np.save(out, film.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_pass_equals_single_rank(tmp_path):
    out = str(tmp_path / "film.npy")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    w, rgba = helpers.emu_pt_render(_assemble())
    exp = np.concatenate([w.reshape(-1), rgba.reshape(-1)])
    assert np.allclose(got, exp, rtol=1e-5, atol=1e-6)
    # and the samples of the two shards are disjoint and complete: weights add up exactly for a box filter
    assert abs(float(got[: w.size].sum()) - float(w.sum())) < 1e-2
