"""Host-side mirror of the reference's LM criterion (misc/utils.py:62-75) for the loss the eval loop reports
(eval_utils.py:139-152).  `forward` keeps the reference's signature on the full [N, S, V+1] log-prob tensor
(plain tensor glue); `from_target_logp` is the form the B200 path uses: the gather already happened inside the
teacher-forced pass (echr_teacher_forced_nll), so only the masked mean is left."""
import torch.nn as nn


class LanguageModelCriterion(nn.Module):
    def __init__(self):
        super(LanguageModelCriterion, self).__init__()

    def forward(self, input, target, mask):
        steps = input.size(1)
        target = target[:, :steps].to(input.device).long()
        mask = mask[:, :steps].to(device=input.device, dtype=input.dtype)
        picked = input.gather(2, target.unsqueeze(2)).squeeze(2)
        return self.from_target_logp(picked, mask)

    @staticmethod
    def from_target_logp(target_logp, mask):
        """target_logp [N, steps] = log p(label_{t+1}); mask [N, >= steps].  -sum(lp*mask) / (sum(mask) + 1e-6)."""
        mask = mask[:, : target_logp.size(1)].to(device=target_logp.device, dtype=target_logp.dtype)
        return -(target_logp * mask).sum() / (mask.sum() + 1e-6)
