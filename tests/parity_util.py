"""Shared parity bars of the GPU tests: 1e-5 relative (north_star) with an absolute floor scaled to the
magnitude of the terms that were summed (the FM second-order term cancels)."""
import numpy as np

RTOL = 1e-5


def close(got, exp, scale):
    """|got - exp| <= RTOL * (|exp| + scale): scale = magnitude of the terms summed"""
    got, exp = np.asarray(got, np.float64), np.asarray(exp, np.float64)
    err = np.abs(got - exp)
    tol = RTOL * (np.abs(exp) + scale)
    tol = np.broadcast_to(tol, err.shape)
    assert (err <= tol).all(), "max err %g, tol there %g" % (err.max(), tol.flat[np.argmax(err - tol)])


def pred_scale(f):
    """per-example magnitude of the terms pred sums (f: a RefModel forward / train_step dict):
    sum |x| |b| + 0.5 sum_k (sum |x v_k|)^2 -- the second-order term cancels, so a 1e-5 relative bar on the
    result alone would be a bar on nothing; this is the floor test_scorer_and_grad_vs_oracle uses too"""
    r = np.abs(f["lp"][f["local"]]).astype(np.float64)
    absx = np.abs(f["vals"]).astype(np.float64)
    ps = f["poses"].astype(np.int64)
    lin = np.concatenate([[0.0], np.cumsum(absx * r[:, 0])])
    fac = np.concatenate([np.zeros((1, r.shape[1] - 1)), np.cumsum(absx[:, None] * r[:, 1:], axis=0)])
    return (lin[ps[1:]] - lin[ps[:-1]]) + 0.5 * ((fac[ps[1:]] - fac[ps[:-1]]) ** 2).sum(1)


def assert_model_state(m, ref, k, lr, steps):
    """trained table and Adagrad accumulator against the oracle model at RTOL; a weight moves by at most lr per
    step (|g| / sqrt(acc) <= 1), which is the scale of what was summed into it"""
    close(m.table[:, :k + 1].cpu().numpy(), ref.flat_table(), lr * steps)
    close(m.accum[:, :k + 1].cpu().numpy(), ref.flat_acc(), 0.0)
