"""Debug: per-member / per-column error pattern of predict vs the oracle for one env."""
import sys, numpy as np
sys.path.insert(0, ".")
from tests.helpers import make_params, make_inputs, orc, out_scale
from tests.test_gpu_predict import _cuda_model
env = sys.argv[1] if len(sys.argv) > 1 else "halfcheetah"
hidden = (200, 200, 200, 200)
p = make_params(env, hidden)
obs, act = make_inputs(env, 2000)
x = np.concatenate([obs, act], -1)
m = _cuda_model(p, hidden, "MSPE")
mean, var = m.predict(x, factored=True)
rm, rv = orc.predict(p, x.astype(np.float64), factored=True)
err = np.abs(mean - rm) / out_scale(p)
np.set_printoptions(precision=3, linewidth=200, suppress=True)
print("env", env, "in_dim", p.in_dim, "out_dim", p.out_dim)
print("per-member max err/s_out:", err.max((1, 2)))
print("per-column max err/s_out:", err.max((0, 1)))
print("per-row-block(128) max:", np.array([err[:, i:i + 128].max() for i in range(0, 2000, 128)]))
lv = np.abs(np.log(var) - np.log(rv))
print("logvar per-column max err:", lv.max((0, 1)))
mean2, _ = m.predict(x, factored=True)
print("repeatable:", np.array_equal(mean, mean2))
