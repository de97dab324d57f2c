"""CPU restatement of the K3b training primitives (TEST INFRASTRUCTURE ONLY -- never imported by the product).

numpy versions of the primitives `uc-nerf_b200/train_path.py` drives (nn.Linear forward / grad-input /
grad-weight, FiLM multiply + ReLU, h * feats_bias, uncertainty-weighted blend: network/models.py:138-184 and its
autograd graph).  With these plugged into `mlp_forward_train` / `mlp_backward` the host-side layer sequence is
checked on CPU against the reference's own autograd gradients (tests/golden/grad_*.npz, made by
tests/golden/make_golden_grad.py); on the GPU the same sequence runs on the CUDA primitives and is compared with
both.  Arguments are torch CPU tensors (possibly strided views); every op works in place through `.numpy()`.
"""
import numpy as np


def _n(t):
    return t.numpy()


def _embed(x, L):
    """network/models.py:50-54: [x | sin(2^l x) l<L | cos(2^l x) l<L]."""
    out = [x]
    out += [np.sin(x * np.float32(2.0 ** l)) for l in range(L)]
    out += [np.cos(x * np.float32(2.0 ** l)) for l in range(L)]
    return np.concatenate(out, -1).astype(np.float32)


class OracleOps:
    def linear_fwd(self, X, W, b, Y, accumulate=False, relu=False):
        y = _n(X).astype(np.float64) @ _n(W).astype(np.float64).T
        if b is not None:
            y = y + _n(b).astype(np.float64)
        if accumulate:
            y = y + _n(Y)
        if relu:
            y = np.maximum(y, 0.0)
        _n(Y)[...] = y.astype(np.float32)

    def linear_bwd_input(self, dY, W, dX, accumulate=False):
        g = _n(dY).astype(np.float64) @ _n(W).astype(np.float64)
        if accumulate:
            g = g + _n(dX)
        _n(dX)[...] = g.astype(np.float32)

    def linear_bwd_weight(self, dY, X, dW, db):
        g = _n(dY).astype(np.float64)
        _n(dW)[...] += (g.T @ _n(X).astype(np.float64)).astype(np.float32)
        if db is not None:
            _n(db)[...] += g.sum(0).astype(np.float32)

    def prep(self, ndc, angle, S, PH, FTV):
        _n(PH)[:, :63] = _embed(_n(ndc), 10)
        _n(FTV)[:, 128:155] = np.repeat(_embed(_n(angle), 4), S, axis=0)

    def film_relu_fwd(self, PRE, MOD, H):
        _n(H)[...] = np.maximum(_n(PRE) * _n(MOD), np.float32(0))

    def film_relu_bwd(self, dH, PRE, MOD, dPRE, dMOD, accumulate, db):
        g = np.where(_n(PRE) * _n(MOD) > 0, _n(dH), np.float32(0))
        _n(dPRE)[...] = g * _n(MOD)
        dm = g * _n(PRE)
        _n(dMOD)[...] = _n(dMOD) + dm if accumulate else dm
        if db is not None:
            _n(db)[...] += _n(dPRE).astype(np.float64).sum(0).astype(np.float32)

    def mul_fwd(self, A, B, OUT):
        _n(OUT)[...] = _n(A) * _n(B)

    def mul_bwd(self, dOUT, A, B, dA, accumulate, dB):
        v = _n(dOUT) * _n(B)
        _n(dA)[...] = _n(dA) + v if accumulate else v
        _n(dB)[...] = _n(dOUT) * _n(A)

    def relu_bwd(self, dY, Y, dX):
        _n(dX)[...] = np.where(_n(Y) > 0, _n(dY), np.float32(0))

    def blend_fwd(self, BASE4, ADAPT4, conf, raw):
        u = 1.0 - _n(conf).astype(np.float64)                                   # models.py:149
        b, a = _n(BASE4).astype(np.float64), _n(ADAPT4).astype(np.float64)
        rgb = 1.0 / (1.0 + np.exp(-(b[:, :3] * (1 - u) + a[:, :3] * u)))        # :177
        s = np.maximum(a[:, 3:] * (1 - u) + b[:, 3:] * u, 0.0)                  # :178
        _n(raw)[...] = np.concatenate([rgb, s], -1).astype(np.float32)

    def blend_bwd(self, d_raw, raw, BASE4, ADAPT4, conf, dBASE4, dADAPT4, d_conf):
        u = 1.0 - _n(conf).astype(np.float64)
        cf = 1.0 - u
        g, o = _n(d_raw).astype(np.float64), _n(raw).astype(np.float64)
        b, a = _n(BASE4).astype(np.float64), _n(ADAPT4).astype(np.float64)
        dx = g[:, :3] * o[:, :3] * (1.0 - o[:, :3])
        dw = np.where(o[:, 3:] > 0, g[:, 3:], 0.0)
        _n(dBASE4)[...] = np.concatenate([dx * cf, dw * u], -1).astype(np.float32)
        _n(dADAPT4)[...] = np.concatenate([dx * u, dw * cf], -1).astype(np.float32)
        du = (dx * (a[:, :3] - b[:, :3])).sum(-1, keepdims=True) + dw * (b[:, 3:] - a[:, 3:])
        _n(d_conf)[...] = (-du).astype(np.float32)
