"""`foldingdiff.plotting` names the GeoBPE drivers use (bin/encode.py:6 star-import, bin/induce.py:17).  The codebook
statistics are computed; everything that draws is a no-op behind the reference's function names (no graphics stack on
this box: "unavailable on this box")."""
import torch


def get_codebook_utility(input_ids, vocab_size, eps=1e-8):
    """plotting.py:78-95: perplexity / entropy / fraction of used ids of a token-id stream (one bincount)"""
    input_ids = torch.as_tensor(input_ids)
    if torch.cuda.is_available() and input_ids.numel() > (1 << 16):
        input_ids = input_ids.cuda(non_blocking=True)
    counts = torch.bincount(input_ids, minlength=vocab_size)
    probs = counts / torch.sum(counts)
    entropy = -torch.sum(probs * torch.log(probs + eps), dim=-1)
    perplexity = torch.exp(entropy)
    out = {
        "perplexity": perplexity,
        "perplexity_normalized": perplexity / vocab_size,
        "entropy": entropy,
        "entropy_normalized": entropy / vocab_size,
        "use_ratio": torch.count_nonzero(counts) / len(counts),
    }
    return {k: v.item() for k, v in out.items()}


def _no_graphics(*args, **kwargs):
    return None


plot = plot_backbone = plot_times = save_histogram = save_histogram_equal_counts = sorted_bar_plot = _no_graphics
plot_joint_kde = plot_val_dists_at_t = plot_losses = plot_consecutive_heatmap = plot_feature_importance = _no_graphics
