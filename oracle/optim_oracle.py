"""TEST INFRASTRUCTURE ONLY (imported by tests/ and bench.py's baseline legs, never by the product path).

CPU restatement (numpy, float32 arithmetic) of the step after the backward:
  * ``adam_step``: torch.optim.Adam as the reference configures it (scene/gaussian_model.py:452: no weight decay, no
    amsgrad), following torch/optim/adam.py ``_single_tensor_adam``: lerp, mul/addcmul, bias corrections as Python
    floats, ``denom = sqrt(v) / sqrt(bc2) + eps``, ``addcdiv``.  PINNED: tests/test_oracle_golden.py runs it against
    ``torch.optim.Adam`` itself (the optimizer the reference calls, train.py:192-194) on CPU.
  * ``densification_stats``: train.py:166-167 and ``GaussianModel.add_densification_stats``
    (scene/gaussian_model.py:756-760).  scene/gaussian_model.py cannot be imported here (tinycudann, pytorch3d absent),
    so this one is pinned against the same torch index expressions restated line by line in the test.
"""
import math

import numpy as np


def adam_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-15):
    """One update, in place on float32 arrays; ``step`` is 1-based."""
    f = np.float32
    m += f(1.0 - beta1) * (g - m)
    v *= f(beta2)
    v += (f(1.0 - beta2) * g) * g
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    step_size = f(lr / bc1)
    denom = np.sqrt(v) * (f(1.0) / f(math.sqrt(bc2))) + f(eps)
    p += (-step_size) * (m / denom)


def densification_stats(vgrad, radii, max_radii2D, accum, denom, accum_max):
    vis = radii > 0
    max_radii2D[vis] = np.maximum(max_radii2D[vis], radii[vis].astype(np.float32))
    g = np.sqrt(vgrad[vis, 0] ** 2 + vgrad[vis, 1] ** 2).astype(np.float32)[:, None]
    accum[vis] += g
    denom[vis] += 1
    accum_max[vis] = np.where(g > accum_max[vis], g, accum_max[vis])
