"""Validation Dice -- drop-in for the reference's unet/evaluate.py:10-43.

The forward runs on the sm_100a kernels; argmax / sigmoid>0.5, the one-hot masks and the per-image Dice
sums are ONE fused kernel producing exact integer counts (bit-exact masks and counts from identical
logits), from which the Dice ratio is evaluated in fp32 like util/dice_score.py:12-17.
"""
import torch

from .. import ops


def dice_from_counts(counts: torch.Tensor, epsilon: float = 1e-6) -> torch.Tensor:
    """counts [n, k, 3] int64 (inter, pred, true) -> mean over (image, class) of the Dice ratio."""
    inter = (2 * counts[..., 0]).float()
    sets = (counts[..., 1] + counts[..., 2]).float()
    sets = torch.where(sets == 0, inter, sets)
    return ((inter + epsilon) / (sets + epsilon)).mean()


def predict_counts(net, image: torch.Tensor, mask_true: torch.Tensor):
    logits = net(image)
    n, c = logits.shape[0], logits.shape[1]
    labels = mask_true.reshape(n, logits.shape[2], logits.shape[3])
    return ops.eval_counts(logits.float().contiguous(), labels.contiguous())


def evaluate(net, dataloader, device, amp, sharded=False):
    """``sharded=True`` (an addition; default = the reference's behaviour): under ``parallel.init_data_parallel`` rank r
    scores the batches i with i % world == r and the ranks exchange (sum of scores, number of batches) -- every rank
    returns the same value as one process scoring all batches."""
    from .. import parallel
    world, rank = (parallel.world_size(), parallel.rank()) if sharded else (1, 0)
    net.eval()
    num_val_batches = len(dataloader)
    dice_score = 0
    scored = 0
    # `amp` only selected fp16 autocast for the reference's cuDNN forward (evaluate.py:16); our forward
    # always computes in the engine's precision (bf16 with fp32 accumulation by default).
    for i, batch in enumerate(dataloader):
        if i % world != rank:
            continue
        scored += 1
        image, mask_true = batch['image'], batch['mask']
        image = image.to(device=device, dtype=torch.float32)
        mask_true = mask_true.to(device=device, dtype=torch.long)
        if net.n_classes == 1:
            assert mask_true.min() >= 0 and mask_true.max() <= 1, 'True mask indices should be in [0, 1]'
        else:
            assert mask_true.min() >= 0 and mask_true.max() < net.n_classes, \
                'True mask indices should be in [0, n_classes['
        counts, _ = predict_counts(net, image, mask_true)
        dice_score += dice_from_counts(counts)
    net.train()
    if sharded and world > 1:
        return parallel.mean_over_ranks(dice_score, scored)
    return dice_score / max(num_val_batches, 1)
