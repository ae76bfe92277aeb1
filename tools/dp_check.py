"""torchrun --nproc-per-node N tools/dp_check.py : N-rank (sync-BN + global Dice) == 1-rank at the same global batch."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import semantic_segmentation_b200 as S
from oracle import genseg_oracle as O
from semantic_segmentation_b200 import parallel
from semantic_segmentation_b200.unet import UNet
from semantic_segmentation_b200.util.dice_score import seg_loss

def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
for prec in ("fp32", "bf16"):
    S.set_precision(prec)
    st = O.seeded_state(O.unet_state_spec(1, 2, True), 1)
    GB = int(os.environ.get("DP_GB", "8"))
    x, y = O.seeded_inputs(2, GB, 1, 64, 64, 2)
    def run(xs, ys):
        net = UNet(1, 2, True); net.load_state_dict(st); net = net.to(dev).train()
        loss = seg_loss(net(xs.to(dev)), ys.to(dev)); loss.backward()
        return loss.detach(), {k: p.grad.clone() for k, p in net.named_parameters()}, net
    parallel.shutdown()
    ref_loss, ref_g, _ = run(x, y)                         # single-rank, full global batch
    world = parallel.init_data_parallel(sync_bn=True, overlap_grad_allreduce=os.environ.get("DP_OVERLAP", "0") == "1")
    lo, hi = parallel.shard_batch(GB)
    loss, g, net = run(x[lo:hi], y[lo:hi])
    parallel.allreduce_gradients(net.parameters())
    g = {k: p.grad for k, p in net.named_parameters()}
    errs = sorted(((rel(g[k], ref_g[k]), k) for k in g), reverse=True)   # global-mean loss -> summed shard grads == full grads
    worst = errs[0][0]
    if parallel.rank() == 0:
        print("world %d %s: loss %.7f ref %.7f  worst grad rel %.3e" % (world, prec, loss.item(), ref_loss.item(), worst), flush=True)
        print("   worst keys: " + ", ".join("%s %.1e" % (k, e) for e, k in errs[:5]) + " | median %.1e" % errs[len(errs) // 2][0], flush=True)
dist.barrier()
