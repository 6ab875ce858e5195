"""Independent second oracle -- TEST INFRASTRUCTURE ONLY.

The dense closed form the reference's own test uses as its expected value
(test/fm_scorer_op_test.py:49-63; stated a second time in
tffm/fm_model.py:246-250), written with torch float64 ops + autograd.  It
shares no code with oracle/fm_oracle.c, so agreement of {C oracle, this,
CUDA kernel} means something.
"""
import torch


def dense_score(fids, params, vals, poses, factor_lambda, bias_lambda):
    """fids/vals flat CSR (like the op); params [R, K+1].  Returns pred[B], reg."""
    params = params.double()
    b = len(poses) - 1
    width = max(int(poses[i + 1] - poses[i]) for i in range(b)) if b else 0
    idx = torch.zeros(b, max(width, 1), dtype=torch.long)
    x = torch.zeros(b, max(width, 1), dtype=torch.float64)
    mask = torch.zeros(b, max(width, 1), dtype=torch.float64)
    for i in range(b):
        n = int(poses[i + 1] - poses[i])
        idx[i, :n] = torch.as_tensor(fids[poses[i]:poses[i + 1]], dtype=torch.long)
        x[i, :n] = torch.as_tensor(vals[poses[i]:poses[i + 1]], dtype=torch.float64)
        mask[i, :n] = 1.0
    factors = params[:, 1:][idx] * mask[..., None]          # [B, W, K]
    biases = params[:, 0][idx] * mask                        # [B, W]
    fx = factors * x[..., None]
    factor_sum = fx.sum(1)
    pred = 0.5 * (factor_sum * factor_sum).sum(1) - 0.5 * (fx * fx).sum((1, 2)) + (biases * x).sum(1)
    reg = 0.5 * (factor_lambda * (factors * factors).sum() + bias_lambda * (biases * biases).sum())
    return pred, reg


def dense_cost_and_grad(fids, params, vals, poses, factor_lambda, bias_lambda, labels):
    """cost = sum((label - pred)^2) + reg and d cost / d params, as in
    test/fm_scorer_op_test.py:60-63."""
    p = params.double().clone().requires_grad_(True)
    pred, reg = dense_score(fids, p, vals, poses, factor_lambda, bias_lambda)
    cost = ((torch.as_tensor(labels, dtype=torch.float64) - pred) ** 2).sum() + reg
    cost.backward()
    return pred.detach(), reg.detach(), p.grad.detach()
