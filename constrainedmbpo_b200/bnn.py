"""B200-native stand-in for ``mbpo.models.bnn.BNN`` / ``mbpo.models.fc.FC`` (forward path only).

Same construction protocol and the same attributes the reference's callers touch
(SURVEY §8b): ``BNN(params)``, ``add(FC(...))``, ``finalize(...)``, ``predict(inputs, factored,
inc_var)``, ``random_inds``, ``elite_inds``, ``num_nets``, ``num_elites``, ``name``,
``is_probabilistic``, ``scaler_in/out.cached_mu/cached_var``, ``save``.  Weights live in CUDA
tensors owned by PyTorch; the arithmetic runs in libcmbpo_b200.so.  Training (``train``) stays
with the reference: its result is handed over with ``load_state`` / ``load_mat``.

All citations are relative to the reference repository root.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _lib


class FC:
    """Layer descriptor with the interface of mbpo/models/fc.py:11-305 that BNN uses."""

    def __init__(self, output_dim, input_dim=None, activation=None, weight_decay=None, ensemble_size=1):
        if activation not in _lib.ACT_IDS and activation != "softmax":
            raise ValueError(f"unknown activation {activation!r} (fc.py:14-21)")
        self.input_dim, self.output_dim = input_dim, output_dim
        self.activation = activation
        self.weight_decay = weight_decay
        self.ensemble_size = ensemble_size

    def __repr__(self):  # fc.py:47-51, the line format of the .nns structure file
        return "FC(output_dim={!r}, input_dim={!r}, activation={!r}, weight_decay={!r}, ensemble_size={!r})".format(
            self.output_dim, self.input_dim, self.activation, self.weight_decay, self.ensemble_size)

    def copy(self):
        return FC(self.output_dim, self.input_dim, self.activation, self.weight_decay, self.ensemble_size)

    def get_input_dim(self): return self.input_dim
    def set_input_dim(self, d): self.input_dim = d
    def get_output_dim(self): return self.output_dim
    def set_output_dim(self, d): self.output_dim = d
    def get_activation(self, as_func=True): return self.activation
    def set_activation(self, a): self.activation = a
    def unset_activation(self): self.activation = None
    def get_ensemble_size(self): return self.ensemble_size
    def set_ensemble_size(self, n): self.ensemble_size = n


def parse_nns(path):
    """Layer list of a ``.nns`` structure file (one ``repr(FC)`` per line, written by ``BNN.save``, bnn.py:811-819;
    read upstream by ``BNN._load_structure``, bnn.py:829-846).  The last line describes the head BEFORE
    ``finalize`` (single-width output for a probabilistic model, its activation restored)."""
    layers = []
    with open(path, "r") as f:
        for line in f:
            line = line.strip()
            if not line:
                continue
            if not (line.startswith("FC(") and line.endswith(")")):
                raise ValueError(f"{path}: not an FC(...) structure line: {line!r}")
            kw = dict(av.split("=", 1) for av in line[3:-1].split(", "))
            act = kw["activation"]
            layers.append(FC(int(kw["output_dim"]), input_dim=int(kw["input_dim"]),
                             activation=None if act == "None" else act.strip("'\""),
                             weight_decay=None if kw["weight_decay"] == "None" else float(kw["weight_decay"]),
                             ensemble_size=int(kw["ensemble_size"])))
    if not layers:
        raise ValueError(f"{path}: empty structure file")
    return layers


class TensorStandardScaler:
    """State holder with the attributes of mbpo/models/utils.py:90-239 callers read
    (``cached_mu``, ``cached_var`` at cmbpo.py:378-379).  ``var`` is a variance although the
    TF variable is named ``_std`` (utils.py:112-115).  ``fit`` keeps the running-moment update
    of utils.py:120-139, :228-239 (host NumPy; it is not on the rollout path)."""

    def __init__(self, x_dim, sc_factor=1, name="Scaler"):
        self.fitted = False
        self.name = name
        self.cached_count, self.cached_mu, self.cached_var = 0, np.zeros([1, x_dim]), np.ones([1, x_dim])
        self.sc_factor = sc_factor
        self._on_change = None

    def set_state(self, mu, var, count=None):
        self.cached_mu = np.asarray(mu, dtype=np.float64).reshape(1, -1)
        self.cached_var = np.asarray(var, dtype=np.float64).reshape(1, -1)
        if count is not None:
            self.cached_count = count
        self.fitted = True
        if self._on_change:
            self._on_change()

    def fit(self, data):
        batch_count = data.shape[0]
        batch_mu = np.mean(data, axis=0, keepdims=True)
        batch_var = np.var(data, axis=0, keepdims=True)
        delta = batch_mu - self.cached_mu
        tot = self.cached_count + batch_count
        new_mean = self.cached_mu + delta * batch_count / tot
        m2 = self.cached_var * self.cached_count + batch_var * batch_count + \
            np.square(delta) * self.cached_count * batch_count / tot
        self.set_state(new_mean, m2 / tot, tot)


def _truncated_normal_(t: torch.Tensor, std: float, gen: torch.Generator):
    """tf.truncated_normal_initializer(stddev) (fc.py:157): redraw outside 2 sigma."""
    torch.nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=gen)
    return t


class BNN:
    """Forward-only ensemble with the surface of mbpo/models/bnn.py:31-903."""

    def __init__(self, params):
        if params.get("name") is None:
            raise ValueError("Must provide name.")  # utils.py:9-13 via bnn.py:49
        self.name = params["name"]
        self.model_dir = params.get("model_dir", None)
        self._sess = params.get("sess", None)  # accepted for signature compatibility, unused
        self.loss_type = params.get("loss", "NLL")
        self.logit_bias_std = params.get("logit_bias_std", 1.0)
        self.finalized = False
        self.layers = []
        self.max_logvar_param = params.get("max_logvar", .5)
        self.min_logvar_param = params.get("min_logvar", -6)
        self.end_act, self.end_act_name = None, None
        self.use_scaler_in = params.get("use_scaler_in", False)
        self.use_scaler_out = params.get("use_scaler_out", False)
        self.sc_factor = params.get("sc_factor", 1)
        self.scaler_in = None
        self.scaler_out = None
        self.num_nets = params.get("num_networks", 1)
        self.num_elites = params["num_elites"]
        # bnn.py:98-102: random elites (duplicates possible, quirk Q7) until the first train()
        self._model_inds = np.random.randint(self.num_nets, size=self.num_elites)
        self.model_loaded = False
        self.device = torch.device(params.get("device", "cuda:0"))
        self.seed = params.get("seed", None)
        # tensor-core operand type of the forward kernels: 'bf16' (default; rtol 1e-2 vs the reference's fp32) or 'tf32'
        # (fp32 operand images, kind::tf32 MMAs, exact activations; rtol 1e-3, about half the throughput)
        self.operand_dtype = params.get("operand_dtype", "bf16")
        if self.operand_dtype not in _lib.DTYPE_IDS:
            raise ValueError(f"operand_dtype must be 'bf16' or 'tf32', got {self.operand_dtype!r}")
        self._handle = None
        self._ws = None
        self.weights, self.biases = [], []
        self.max_logvar = self.min_logvar = None      # NLL models carry them as trained variables (bnn.py:208-212)
        if params.get("load_model", False):           # bnn.py:90-94
            if self.model_dir is None:
                raise ValueError("Cannot load model without providing model directory.")
            self.layers = parse_nns(os.path.join(self.model_dir, "%s.nns" % self.name))
            self.num_nets, self.model_loaded = self.layers[0].get_ensemble_size(), True
            self._model_inds = np.random.randint(self.num_nets, size=self.num_elites)   # bnn.py:96-97

    # ---- properties of the reference ------------------------------------------------------
    @property
    def is_probabilistic(self):  # bnn.py:110-112
        return True if "NLL" in self.loss_type or "MSPE" in self.loss_type else False

    @property
    def is_tf_model(self):
        return False

    @property
    def sess(self):
        return self._sess

    @property
    def elite_inds(self):  # bnn.py:433-435
        return self._model_inds

    def random_inds(self, batch_size):  # bnn.py:437-439 (global NumPy RNG, like the reference)
        return np.random.choice(self._model_inds, size=batch_size)

    # ---- structure ------------------------------------------------------------------------
    def add(self, layer):  # bnn.py:126-145
        if self.finalized:
            raise RuntimeError("Cannot modify network structure after finalizing.")
        if len(self.layers) == 0 and layer.get_input_dim() is None:
            raise ValueError("Must set input dimension for the first layer.")
        if self.model_loaded:
            raise RuntimeError("Cannot add layers to a loaded model.")
        layer.set_ensemble_size(self.num_nets)
        if len(self.layers) > 0:
            layer.set_input_dim(self.layers[-1].get_output_dim())
        self.layers.append(layer.copy())

    def pop(self):  # bnn.py:147-159
        if len(self.layers) == 0:
            raise RuntimeError("Network is empty.")
        if self.finalized:
            raise RuntimeError("Cannot modify network structure after finalizing.")
        return self.layers.pop()

    def finalize(self, optimizer=None, optimizer_args=None, weighted=False, *args, **kwargs):
        """bnn.py:161-380 minus the training graph: double the head of a probabilistic model
        (:190-191), strip the last activation (:194-196), create scalers (:201-207) and
        variables (fc.py:154-185), then build the native handle."""
        if len(self.layers) == 0:
            raise RuntimeError("Cannot finalize an empty network.")
        if self.finalized:
            raise RuntimeError("Can only finalize a network once.")
        if self.is_probabilistic:
            self.layers[-1].set_output_dim(2 * self.layers[-1].get_output_dim())
        self.end_act_name = self.layers[-1].get_activation(as_func=False)
        self.end_act = self.end_act_name
        self.layers[-1].unset_activation()
        if self.end_act_name is not None:
            raise _lib.CmbpoError("output_activation is not supported on the B200 rollout path "
                                  f"(got {self.end_act_name!r}); the dynamics / cost / critic models use None")
        in_dim = self.layers[0].get_input_dim()
        head = self.layers[-1].get_output_dim()
        self.out_dim = head // 2 if self.is_probabilistic else head
        if self.use_scaler_in:
            self.scaler_in = TensorStandardScaler(in_dim, sc_factor=self.sc_factor, name="scaler_in")
        if self.use_scaler_out:
            self.scaler_out = TensorStandardScaler(self.out_dim, sc_factor=self.sc_factor, name="scaler_out")

        lib = _lib.require_device(self.device.index or 0)
        cfg = _lib.cmbpo_cfg()
        cfg.num_members, cfg.num_layers = self.num_nets, len(self.layers)
        cfg.in_dim, cfg.out_dim = in_dim, self.out_dim
        cfg.probabilistic = 1 if self.is_probabilistic else 0
        for i, layer in enumerate(self.layers):
            cfg.layer_width[i] = layer.get_output_dim()
            if layer.get_activation() == "softmax":
                raise _lib.CmbpoError("softmax layers are not supported on the B200 rollout path")
            cfg.layer_act[i] = _lib.ACT_IDS[layer.get_activation()]
        cfg.operand_dtype = _lib.DTYPE_IDS[self.operand_dtype]
        h = C.c_void_p()
        _lib.check(lib.cmbpo_create(C.byref(cfg), self.device.index or 0, C.byref(h)))
        self._handle = h
        self._lib = lib

        gen = torch.Generator(device="cpu")
        gen.manual_seed(int(self.seed) if self.seed is not None else int(np.random.randint(2 ** 31 - 1)))
        self.weights, self.biases = [], []
        for i, layer in enumerate(self.layers):
            k, n = layer.get_input_dim(), layer.get_output_dim()
            std = 1.0 / (2.0 * np.sqrt(k))  # fc.py:157
            if i == len(self.layers) - 1:
                std *= self.logit_bias_std  # bnn.py:216-217
            w = _truncated_normal_(torch.empty(self.num_nets, k, n, dtype=torch.float32), std, gen)
            self.weights.append(w.to(self.device))
            self.biases.append(torch.zeros(self.num_nets, 1, n, dtype=torch.float32, device=self.device))
        for s in (self.scaler_in, self.scaler_out):
            if s is not None:
                s._on_change = self._push_scalers
        if self.loss_type == "NLL":    # bnn.py:208-212
            self.max_logvar = np.ones([1, self.out_dim], np.float32) * self.max_logvar_param
            self.min_logvar = np.ones([1, self.out_dim], np.float32) * self.min_logvar_param
        self.finalized = True
        self._push_weights()
        self._push_scalers()
        self._push_elites()
        if self.model_loaded:          # bnn.py:374-379
            self.load_mat(os.path.join(self.model_dir, "%s.mat" % self.name))

    # ---- state hand-over (after the reference trained, bnn.py:467-666) ----------------------
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _push_weights(self):
        for i, (w, b) in enumerate(zip(self.weights, self.biases)):
            _lib.check(self._lib.cmbpo_set_weights(self._handle, i, C.c_void_p(w.data_ptr()),
                                                   C.c_void_p(b.data_ptr()), self._stream()))

    def _push_scalers(self):
        if not self.finalized:
            return
        keep = []

        def ptr(a):
            if a is None:
                return None
            arr = np.ascontiguousarray(np.asarray(a, dtype=np.float32).reshape(-1))
            keep.append(arr)
            return arr.ctypes.data_as(C.c_void_p)
        si, so = self.scaler_in, self.scaler_out
        _lib.check(self._lib.cmbpo_set_scalers(
            self._handle,
            ptr(si.cached_mu if si else None), ptr(si.cached_var if si else None),
            ptr(so.cached_mu if so else None), ptr(so.cached_var if so else None),
            C.c_float(float(self.sc_factor)), self._stream()))

    def _push_elites(self):
        inds = np.ascontiguousarray(np.asarray(self._model_inds, dtype=np.int32))
        _lib.check(self._lib.cmbpo_set_elites(self._handle, inds.ctypes.data_as(C.POINTER(C.c_int32)),
                                              len(inds), self._stream()))

    def set_elites(self, inds):
        """What ``_end_train`` does with the sorted holdout losses (bnn.py:428-431)."""
        self._model_inds = list(inds)
        self._push_elites()

    def load_state(self, weights, biases, scaler_in=None, scaler_out=None, elites=None):
        """Take over trained parameters: lists of ``[E,in,out]`` / ``[E,1,out]`` arrays, scaler
        ``(mu, var)`` pairs, elite list.  Needed after every ``train()``."""
        assert len(weights) == len(self.layers) and len(biases) == len(self.layers)
        for i, (w, b) in enumerate(zip(weights, biases)):
            w = (w if torch.is_tensor(w) else torch.as_tensor(np.asarray(w))).to(torch.float32)
            b = (b if torch.is_tensor(b) else torch.as_tensor(np.asarray(b))).to(torch.float32)
            if tuple(w.shape) != tuple(self.weights[i].shape) or b.numel() != self.biases[i].numel():
                raise ValueError(f"layer {i}: expected {tuple(self.weights[i].shape)}, got {tuple(w.shape)}")
            self.weights[i] = w.to(self.device).contiguous()
            self.biases[i] = b.reshape(self.biases[i].shape).to(self.device).contiguous()
        self._push_weights()
        if scaler_in is not None and self.scaler_in is not None:
            self.scaler_in.set_state(*scaler_in)
        if scaler_out is not None and self.scaler_out is not None:
            self.scaler_out.set_state(*scaler_out)
        if elites is not None:
            self.set_elites(elites)

    def load_mat(self, path):
        """Ingest the ``.mat`` written by the reference's ``BNN.save`` (bnn.py:821-827): keys '0'..'N' in the order
        nonoptvars + optvars = [mu_in, var_in, mu_out, var_out, W0, b0, ..., W_L, b_L, (max_logvar, min_logvar)]
        (bnn.py:214-227): scaler entries only for the scalers in use, the two log-variance bounds only for 'NLL'."""
        from scipy.io import loadmat
        d = loadmat(path)
        n = len([k for k in d if k.isdigit()])
        vals = [d[str(i)] for i in range(n)]
        L = len(self.layers)
        want = 2 * L + 2 * bool(self.use_scaler_in) + 2 * bool(self.use_scaler_out) + 2 * (self.loss_type == "NLL")
        if n != want:
            raise ValueError(f"{path}: {n} variables, this structure needs {want} "
                             f"(scalers in/out: {self.use_scaler_in}/{self.use_scaler_out}, loss {self.loss_type})")
        i = 0
        sin = sout = None
        if self.use_scaler_in:
            sin = (vals[i], vals[i + 1]); i += 2
        if self.use_scaler_out:
            sout = (vals[i], vals[i + 1]); i += 2
        ws = vals[i:i + 2 * L:2]
        bs = vals[i + 1:i + 2 * L:2]
        for k, b in enumerate(bs):     # savemat squeezes nothing, but tolerate [E, out] biases
            bs[k] = np.asarray(b).reshape(self.num_nets, 1, -1)
        i += 2 * L
        self.load_state(ws, bs, sin, sout)
        if self.loss_type == "NLL":
            self.max_logvar = np.asarray(vals[i], np.float32).reshape(1, -1)
            self.min_logvar = np.asarray(vals[i + 1], np.float32).reshape(1, -1)

    def save(self, savedir, timestep):
        """Same two files as bnn.py:799-827: ``{name}_{timestep}.nns`` + ``.mat``; a reference ``BNN`` constructed
        with ``load_model=True, name='{name}_{timestep}'`` reads them back (and so does this class)."""
        from scipy.io import savemat
        if not self.finalized:
            raise RuntimeError()
        model_dir = self.model_dir if savedir is None else savedir
        with open(os.path.join(model_dir, "{}_{}.nns".format(self.name, timestep)), "w+") as f:
            for layer in self.layers[:-1]:
                f.write("%s\n" % repr(layer))
            last = self.layers[-1].copy()
            last.set_activation(self.end_act_name)
            if self.is_probabilistic:
                last.set_output_dim(last.get_output_dim() // 2)
            f.write("%s\n" % repr(last))
        vals = []
        for s in (self.scaler_in, self.scaler_out):
            if s is not None:
                vals += [np.asarray(s.cached_mu, np.float32), np.asarray(s.cached_var, np.float32)]
        for w, b in zip(self.weights, self.biases):
            vals += [w.cpu().numpy(), b.cpu().numpy()]
        if self.loss_type == "NLL":
            vals += [np.asarray(self.max_logvar, np.float32), np.asarray(self.min_logvar, np.float32)]
        savemat(os.path.join(model_dir, "{}_{}.mat".format(self.name, timestep)),
                {str(i): v for i, v in enumerate(vals)})

    def train(self, *args, **kwargs):
        raise NotImplementedError(
            "training stays with the reference implementation (bnn.py:467-666); hand the trained "
            "state over with BNN.load_state(...) / load_mat(...)")

    def reset(self):
        raise NotImplementedError("re-initialisation belongs to the training side (bnn.py:441-449)")

    # ---- prediction -------------------------------------------------------------------------
    def _workspace(self, n_rows):
        need = self._lib.cmbpo_workspace_bytes(self._handle, n_rows)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(int(need * 1.25) + 1024, dtype=torch.uint8, device=self.device)
        return self._ws

    def predict_device(self, segments, ndim, agg, want_var=True, n_rows_dev=None):
        """Device-level predict on a list of CUDA fp32 segments whose concatenation along the
        last axis is the input (no concat copy is made).  Returns CUDA tensors.  ``n_rows_dev``: optional int32 device
        scalar with the live row count (rollout loop: shapes stay at capacity, rows beyond it are not computed)."""
        if not self.finalized:
            raise RuntimeError("finalize() the model first")
        segs = [s.contiguous() for s in segments]
        n_rows = segs[0].shape[-2]
        E, D = self.num_nets, self.out_dim
        io = _lib.cmbpo_predict_io()
        for i in range(_lib.MAX_SEGMENTS):
            if i < len(segs):
                io.seg_dev[i] = segs[i].data_ptr()
                io.seg_width[i] = segs[i].shape[-1]
            else:
                io.seg_dev[i] = None
                io.seg_width[i] = 0
        io.n_rows, io.ndim, io.agg = n_rows, ndim, agg
        factored = (agg == 0) or ndim == 3
        shape = (E, n_rows, D) if factored else (n_rows, D)
        mean = torch.empty(shape, dtype=torch.float32, device=self.device)
        var = torch.empty(shape, dtype=torch.float32, device=self.device) if want_var else None
        ws = self._workspace(n_rows)
        io.mean_dev, io.var_dev = mean.data_ptr(), (var.data_ptr() if var is not None else None)
        io.workspace_dev, io.workspace_bytes = ws.data_ptr(), ws.numel()
        if n_rows_dev is not None:
            io.n_rows_dev = n_rows_dev.data_ptr()
        _lib.check(self._lib.cmbpo_predict(self._handle, C.byref(io), self._stream()))
        return mean, var

    def predict(self, inputs, factored=False, inc_var=True, *args, **kwargs):
        """bnn.py:671-776; shapes per SURVEY appendix B.  NumPy in -> NumPy out, CUDA tensor
        in -> CUDA tensors out."""
        is_np = not torch.is_tensor(inputs)
        x = torch.as_tensor(np.asarray(inputs, dtype=np.float32)) if is_np else inputs.to(torch.float32)
        x = x.to(self.device, non_blocking=True)
        if x.dim() not in (2, 3) or (x.dim() == 3 and x.shape[0] != self.num_nets) \
                or x.shape[-1] != self.layers[0].get_input_dim():
            raise ValueError("Invalid input dimension.")  # fc.py:110-111
        def out(*ts):      # NumPy in -> NumPy out through pinned host blocks, one synchronisation for all results
            r = _lib.to_host(ts) if is_np else list(ts)
            return r[0] if len(r) == 1 else r
        prob = self.is_probabilistic
        if x.dim() == 2:
            if factored:
                mean, var = self.predict_device([x], 2, 0, want_var=inc_var and prob)
                if inc_var and prob:
                    return list(out(mean, var))  # :713-717 (sess.run returns a list)
                if inc_var:
                    rep = mean.var(dim=0, unbiased=False)[None].repeat(mean.shape[0], 1, 1)  # :718-724
                    return tuple(out(mean, rep))
                return out(mean)
            mean, var = self.predict_device([x], 2, 1, want_var=inc_var)
            if inc_var and prob:
                return list(out(mean, var))  # :731-735
            if inc_var:
                return tuple(out(mean, var))  # :736-743
            return out(mean)
        mean, var = self.predict_device([x], 3, 0, want_var=inc_var and prob)
        if inc_var and prob:
            return list(out(mean, var))  # :750-754
        if inc_var:  # :755-771: variance across members of the flattened 2-D pass
            flat = x.reshape(-1, x.shape[-1])
            m2d, _ = self.predict_device([flat], 2, 0, want_var=False)
            v3d = m2d.var(dim=0, unbiased=False).reshape(x.shape[0], x.shape[1], -1)
            return tuple(out(mean, v3d))
        return out(mean)

    def check_status(self):
        """Synchronise this model's stream and raise if a device watchdog tripped (include/cmbpo_b200.h cmbpo_status)."""
        _lib.check(self._lib.cmbpo_status(self._handle, self._stream()))

    def close(self):
        if self._handle is not None:
            self._lib.cmbpo_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
