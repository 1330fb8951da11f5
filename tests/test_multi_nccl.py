"""N = 2 on real GPUs (NCCL): the product's sharded render (Interface with a comm: samples s mod N per rank, one all-reduce of
the pass film per AA pass, adaptive flags from the all-reduced film) must reproduce the 1-GPU film.  Needs 2 GPUs: skipped on
the 1-GPU box (run with `gpurun --gpus 2 -- python -m pytest tests/test_multi_nccl.py -m gpu`)."""
import os

import numpy as np
import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu
KW = dict(width=160, height=90, spp=8, bounces=5)
EXTRA = dict(AA_passes=3, AA_inc_samples=4, AA_threshold=0.03)


def _render(yi, threshold):
    from libyafaray_b200 import scenes_pt
    scenes_pt.cornell_box(yi, accelerator="yafaray-b200-bvh8", **KW)
    yi.render_params.update(dict(EXTRA, AA_threshold=threshold))
    img = yi.render()
    w, rgba = yi.read_film()
    return img, w, rgba, list(yi.resampled_per_pass)


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from libyafaray_b200 import interface
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    res = {}
    for name, thr in (("all", 0.0), ("adaptive", EXTRA["AA_threshold"])):
        yi = interface.Interface.distributed()
        img, w, rgba, hist = _render(yi, thr)
        res.update({name + "_img": img, name + "_w": w, name + "_rgba": rgba, name + "_hist": np.array(hist)})
        yi.clearAll()
    torch.cuda.synchronize()
    np.savez(out % rank, **res)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_sharded_adaptive_render_equals_the_single_gpu_film(tmp_path):
    import torch.multiprocessing as mp
    from libyafaray_b200 import interface
    out = str(tmp_path / "rank%d.npz")
    mp.spawn(_worker, args=(2, 29500 + (os.getpid() % 2000), out), nprocs=2, join=True)
    r0, r1 = np.load(out % 0), np.load(out % 1)
    # every rank holds the complete film, bit for bit (the all-reduce hands all ranks the same sums)
    for k in r0.files:
        assert np.array_equal(r0[k], r1[k]), k
    # every pass over every pixel: the same samples summed in another order -> fp32 reordering only
    img, w, rgba, hist = _render(interface.Interface(), 0.0)
    assert list(r0["all_hist"]) == hist
    assert np.allclose(r0["all_w"], w, rtol=0, atol=1e-4) and np.allclose(r0["all_rgba"], rgba, rtol=1e-5, atol=1e-4)
    assert helpers.psnr(r0["all_img"], img) >= 100.0
    # adaptive passes: the pixels are flagged from the all-reduced film; a pixel whose noise sits exactly on the threshold may
    # flip with the summation order, so the flagged counts agree closely rather than exactly
    img, w, rgba, hist = _render(interface.Interface(), EXTRA["AA_threshold"])
    assert sum(hist) > 0
    assert all(abs(int(a) - int(b)) <= max(4, 0.01 * b) for a, b in zip(r0["adaptive_hist"], hist)), (list(r0["adaptive_hist"]), hist)
    assert helpers.psnr(r0["adaptive_img"], img) >= 45.0
