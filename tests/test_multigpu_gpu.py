"""2-rank GPU test (needs >= 2 GPUs: run under `gpurun --gpus 2`): tile-sharded frame gathered over peer memory (fused pull +
finalize kernel) and over NCCL == the single-GPU frame, also when converging over several frames; sample-sharded over NCCL."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
import cases, test_wavefront_gpu as tw
from lighthouse2_b200 import core, multigpu, scenes
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
case = tw._scene(); w, h = 200, 100
view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
for mode, gather in (("tiles", "p2p"), ("tiles", "nccl"), ("samples", "nccl")):
    sr = multigpu.ShardedRenderer(rank, world, local, mode=mode, gather=gather)
    cases.load_core(sr.core, dict(case, rays=None))
    sr.SetTarget(w, h, 1); sr.SetPointLight(*tw.LIGHT); sr.SetSky(tw.SKY); sr.Setting("maxPathLength", 4); sr.Setting("jitter", 0)
    assert sr.gather == gather
    sr.Render(view, 1)
    got = sr.ReadAccumulator()
    frame = sr.ReadFrame()
    got3 = None
    if mode == "tiles":  # converge: two more frames on top
        sr.Render(view, 0); sr.Render(view, 0)
        got3, frame3 = sr.ReadAccumulator(), sr.ReadFrame()
    sr.Shutdown()
    if rank == 0:
        ref = tw._core(case, w, h, spp=1 if mode == "tiles" else world)
        ref.Render(view, 1)
        full = ref.ReadAccumulator(); full_frame = ref.ReadFrame()
        if mode == "tiles":
            assert np.array_equal(got.view(np.uint32), full.view(np.uint32)), "tile-sharded gather must be bit-identical"
            assert np.array_equal(frame.view(np.uint32), full_frame.view(np.uint32))
            ref.Render(view, 0); ref.Render(view, 0)
            full3, full_frame3 = ref.ReadAccumulator(), ref.ReadFrame()
            if gather == "p2p":  # every pixel is accumulated by its owner alone, in the single core's order
                assert np.array_equal(got3.view(np.uint32), full3.view(np.uint32)) and np.array_equal(frame3.view(np.uint32), full_frame3.view(np.uint32))
            else:
                assert np.allclose(got3, full3, rtol=1e-6, atol=1e-7) and np.allclose(frame3, full_frame3, rtol=1e-6, atol=1e-7)
        else:
            assert np.allclose(got[..., :3], full[..., :3], rtol=1e-5, atol=1e-6)
        ref.Shutdown()
        print("multi-gpu", mode, gather, "ok")
dist.barrier(); dist.destroy_process_group()
""" % (ROOT, ROOT)


def test_two_rank_frame_equals_single_gpu(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29611", str(script)], capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-3000:])
    assert r.returncode == 0
    assert "multi-gpu tiles p2p ok" in r.stdout and "multi-gpu tiles nccl ok" in r.stdout and "multi-gpu samples nccl ok" in r.stdout
