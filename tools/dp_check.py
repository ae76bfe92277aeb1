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
    # one optimiser step: seg_loss is a GLOBAL-batch mean, so the summed shard gradients are the full-batch gradient and the
    # optimiser must not divide by the world size (grad_scale = 1): N-rank parameters after the step == 1-rank parameters
    from semantic_segmentation_b200.optim import RMSprop
    parallel.shutdown()
    ref_net = UNet(1, 2, True); ref_net.load_state_dict(st); ref_net = ref_net.to(dev).train()
    for k, p in ref_net.named_parameters():
        p.grad = ref_g[k].clone()
    RMSprop(ref_net.parameters(), lr=1e-4, weight_decay=1e-8, momentum=0.9).step()
    opt = RMSprop(net.parameters(), lr=1e-4, weight_decay=1e-8, momentum=0.9)
    opt.grad_scale = 1.0
    opt.step()
    ref_p = dict(ref_net.named_parameters())
    step_err = max(float((p.detach() - ref_p[k].detach()).abs().max()) for k, p in net.named_parameters())
    if int(os.environ.get("RANK", "0")) == 0:
        print("world %d %s: loss %.7f ref %.7f  worst grad rel %.3e  max |param - 1-rank param| after one RMSprop step %.3e (lr 1e-4)"
              % (world, prec, loss.item(), ref_loss.item(), worst, step_err), flush=True)
        print("   worst keys: " + ", ".join("%s %.1e" % (k, e) for e, k in errs[:5]) + " | median %.1e" % errs[len(errs) // 2][0], flush=True)
# rank-count-invariant dropout: the generator's Philox masks are indexed by the GLOBAL element, so a sharded batch draws the
# masks of the one-rank run (sync-BN on): the 2-rank outputs equal the matching rows of the 1-rank output
from semantic_segmentation_b200.models_pix2pix import networks
S.set_precision("fp32")
parallel.shutdown()
torch.manual_seed(77)
G = networks.define_G(1, 1, 64, "unet_256", "batch", True, "normal", 0.02, []).to(dev).train()
networks.upconv_arch = (1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(dev).requires_grad_(True)
GBD = max(4, int(os.environ.get("WORLD_SIZE", "1")))      # one image per rank at 8 ranks
m = torch.randint(0, 2, (GBD, 1, 256, 256), generator=torch.Generator().manual_seed(6)).float()
sd = {k: v.clone() for k, v in G.state_dict().items()}
torch.manual_seed(99); G._program._calls = 0
with torch.no_grad():
    full = G(m.to(dev))
world = parallel.init_data_parallel(sync_bn=True)
G.load_state_dict(sd)
lo, hi = parallel.shard_batch(GBD)
torch.manual_seed(99); G._program._calls = 0
with torch.no_grad():
    part = G(m[lo:hi].to(dev))
e = rel(part, full[lo:hi])
allv = [None] * world
dist.all_gather_object(allv, e)
if int(os.environ.get("RANK", "0")) == 0:
    print("dropout invariance world %d: rel-L2 of the sharded generator output against the one-rank rows: %s" % (world, ", ".join("%.2e" % v for v in allv)), flush=True)
S.set_precision("bf16")
dist.barrier()
