"""CPU restatement of the masked action head, generic/agent.py:190-277.
TEST INFRASTRUCTURE - see oracle/__init__.py.  Pinned by tests/golden/head.npz
(the reference's Agent methods executed verbatim)."""
import torch


def action_and_logprob(logits, action=None, action_masks=None, deterministic=False,
                       generator=None):
    """generic/agent.py:197-231 on precomputed ``logits [B, n]``.
    Returns ``(action, logprob, entropy, masked_logits)``."""
    mask = None
    if action_masks is not None:
        mask = torch.as_tensor(action_masks[:, : logits.shape[1]], dtype=torch.bool)   # :202-205
        logits = torch.where(mask, logits, torch.tensor(-1e20, dtype=logits.dtype))  # :206-207
    logp = logits - logits.logsumexp(dim=-1, keepdim=True)       # Categorical(logits=...) normalisation
    probs = torch.softmax(logits, dim=-1)                        # Categorical.probs
    if action is None:
        if not deterministic:
            action = torch.multinomial(probs, 1, True, generator=generator).squeeze(-1)  # distrib.sample()
        else:
            action = torch.argmax(probs, dim=1)                  # :214
    if mask is None:
        p_log_p = logp.clamp(min=torch.finfo(logp.dtype).min) * probs   # Categorical.entropy()
        entropy = -p_log_p.sum(-1)
    else:
        nlogits = logits - logits.logsumexp(dim=-1, keepdim=True)        # :218
        p_log_p = nlogits * torch.softmax(nlogits, dim=-1)               # :219 logits_to_probs
        p_log_p = torch.where(mask, p_log_p, torch.tensor(0.0, dtype=logits.dtype))
        entropy = -p_log_p.sum(-1)                                       # :223
    logprob = logp.gather(-1, action.long().unsqueeze(-1)).squeeze(-1)   # distrib.log_prob(action)
    return action, logprob, entropy, logits


def action_probs(logits, action_masks):
    """generic/agent.py:238-254 (pretraining): fill -1e10, plain softmax."""
    if action_masks is not None:
        mask = torch.as_tensor(action_masks[:, : logits.shape[1]], dtype=torch.bool)
        logits = torch.where(mask, logits, torch.tensor(-1e10, dtype=logits.dtype))
    return torch.softmax(logits, dim=-1)


def predict(logits, deterministic, action_masks, generator=None):
    """generic/agent.py:256-277 on ``logits [B, n]`` (inference): argmax of the masked
    *logits* (not probs) when deterministic."""
    if action_masks is not None:
        mask = torch.as_tensor(action_masks[:, : logits.shape[1]], dtype=torch.bool)
        logits = torch.where(mask, logits, torch.tensor(-1e20, dtype=logits.dtype))
    if not deterministic:
        return torch.multinomial(torch.softmax(logits, -1), 1, True, generator=generator).squeeze(-1)
    return torch.argmax(logits, dim=-1)
