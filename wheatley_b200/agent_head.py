"""Masked action head - the arithmetic of ``Agent.get_action_and_value`` /
``get_action_probs_and_value`` / ``predict`` (generic/agent.py:190-277) after ``self.gnn(x)``,
on the fused CUDA kernels ``b200gat_mask_head_fwd/bwd``.

``AgentHeadMixin`` gives an agent class the three reference methods with the reference
signatures and return tuples; it expects ``self.gnn``, ``self.value_net``, ``self.action_net``
exactly like the reference ``Agent`` (generic/agent.py:143-147)."""
import torch
from torch.distributions.categorical import Categorical

from .functional import masked_log_softmax, masked_logprob_entropy

class LazyCategorical:
    """``Categorical(logits=logits)`` of generic/agent.py:197, built on first use.  The update loop reads it only
    for the optional uniform-target KL term (alg/ppo.py:694-699); constructing it costs a log-sum-exp and a
    subtraction over ``[B, n]`` plus their autograd nodes every minibatch.  Attribute access (``.probs``,
    ``.logits``, ``.entropy()``, ``.log_prob()``, ``.sample()`` ...) is forwarded to the real distribution.
    ``validate_args=False``: the default argument check reads a device flag back to the host (a sync per call,
    illegal under CUDA-graph capture); the distribution itself is identical."""

    def __init__(self, logits):
        self._logits, self._d = logits, None

    def materialize(self) -> Categorical:
        if self._d is None:
            self._d = Categorical(logits=self._logits, validate_args=False)
        return self._d

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self.materialize(), name)


HUGE_NEG = -1e20          # generic/agent.py:206, :270
HUGE_NEG_PRETRAIN = -1e10  # generic/agent.py:245-247


def action_logprob_entropy(logits, action=None, action_masks=None, deterministic=False):
    """generic/agent.py:197-231 on ``logits [B, n]``: returns ``(action, logprob, entropy)``.

    * masks are cropped to ``logits.shape[1]`` (:202-203) and masked logits are filled with -1e20;
    * ``action is None``: sample from the masked categorical, or first-index argmax when
      ``deterministic`` (:210-214);
    * entropy is ``-sum_mask p log p`` (:218-223) (all entries without a mask, ``Categorical.entropy``)."""
    if action is None:
        logp, entropy, amax = masked_log_softmax(logits, action_masks, HUGE_NEG)
        if deterministic:
            action = amax
        else:
            # Categorical.sample() == torch.multinomial(probs, 1, True): same RNG stream as the
            # reference for identical probabilities
            action = torch.multinomial(logp.detach().exp(), 1, True).squeeze(-1)
        logprob = logp.gather(-1, action.unsqueeze(-1)).squeeze(-1)
        return action, logprob, entropy
    logprob, entropy = masked_logprob_entropy(logits, action, action_masks, HUGE_NEG)
    return action, logprob, entropy


def action_probs(logits, action_masks=None):
    """generic/agent.py:238-254 (pretraining): fill -1e10, plain softmax."""
    logp, _, _ = masked_log_softmax(logits, action_masks, HUGE_NEG_PRETRAIN)
    return logp.exp()


def predict_action(logits, deterministic, action_masks=None):
    """generic/agent.py:256-277: argmax over the masked logits, or a sample."""
    with torch.no_grad():
        logp, _, amax = masked_log_softmax(logits.detach(), action_masks, HUGE_NEG)
        if deterministic:
            return amax
        return torch.multinomial(logp.exp(), 1, True).squeeze(-1)


class AgentHeadMixin:
    def get_value(self, x):
        _, graph_embedding = self.gnn(x)
        return self.value_net(graph_embedding)

    def get_action_and_value(self, x, action=None, action_masks=None, deterministic=False):
        node_features, graph_embedding = self.gnn(x)
        value = self.value_net(graph_embedding)
        logits = self.action_net(node_features).squeeze(-1)
        unmasked_distrib = LazyCategorical(logits)                  # generic/agent.py:197
        action, logprob, entropy = action_logprob_entropy(logits, action, action_masks, deterministic)
        return action, logprob, entropy, value, unmasked_distrib

    def get_action_probs_and_value(self, x, action_masks):
        node_features, graph_feature = self.gnn(x)
        value = self.value_net(graph_feature)
        logits = self.action_net(node_features).squeeze(-1)
        return action_probs(logits, action_masks), value

    def predict(self, observation, deterministic, action_masks):
        with torch.no_grad():
            features, _ = self.gnn(observation)
            logits = self.action_net(features).squeeze(-1)
            return predict_action(logits, deterministic, action_masks)

    def forward(self, observation, action=None, action_masks=None, deterministic=False):
        return self.get_action_and_value(observation, action, action_masks, deterministic)
