"""``construct_model`` with the signature of mbpo/models/constructor.py:9-80."""
from __future__ import annotations

from .bnn import BNN, FC


def construct_model(in_dim,
                    out_dim,
                    name='BNN',
                    hidden_dims=(200, 200, 200),
                    num_networks=7,
                    num_elites=5,
                    loss='NLL',
                    activation='swish',
                    output_activation=None,
                    decay=1e-4,
                    lr=1e-3,
                    lr_decay=None,
                    decay_steps=None,
                    weighted=False,
                    use_scaler_in=False,
                    use_scaler_out=False,
                    sc_factor=1,
                    clip_loss=False,
                    kl_cliprange=0.1,
                    var_corr=False,
                    max_logvar=.5,
                    min_logvar=-6,
                    logit_bias_std=1.0,
                    session=None,
                    device="cuda:0",
                    seed=None,
                    operand_dtype="bf16"):
    """Builds the ensemble the way constructor.py:41-67 does: ``BNN(params)``, one ``FC`` per
    hidden dim (weight decays ``decay/4``, ``decay/2``, ``decay``), ``finalize``.  Optimiser
    arguments (``lr``, ``lr_decay``, ``decay_steps``, ``weighted``, ``clip_loss``, ...) are
    accepted and recorded but unused: training stays with the reference.  ``device`` / ``seed``
    are additions (CUDA device of the handle; seed of the random initial weights), and so is ``operand_dtype``
    ('bf16' | 'tf32': the tensor-core operand type, include/cmbpo_b200.h cmbpo_cfg.operand_dtype)."""
    params = {'name': name,
              'loss': loss,
              'num_networks': num_networks,
              'num_elites': num_elites,
              'sess': session,
              'use_scaler_in': use_scaler_in,
              'use_scaler_out': use_scaler_out,
              'sc_factor': sc_factor,
              'clip_loss': clip_loss,
              'kl_cliprange': kl_cliprange,
              'var_corr': var_corr,
              'max_logvar': max_logvar,
              'min_logvar': min_logvar,
              'logit_bias_std': logit_bias_std,
              'device': device,
              'seed': seed,
              'operand_dtype': operand_dtype,
              }
    model = BNN(params)
    model.add(FC(hidden_dims[0], input_dim=in_dim, activation=activation, weight_decay=decay / 4))
    for hidden_dim in hidden_dims[1:]:
        model.add(FC(hidden_dim, activation=activation, weight_decay=decay / 2))
    model.add(FC(out_dim, activation=output_activation, weight_decay=decay))
    opt_params = {"learning_rate": lr} if lr_decay is None else {"learning_rate": lr,
                                                                 "learning_rate_decay": lr_decay,
                                                                 "decay_steps": decay_steps}
    model.finalize(None, opt_params, weighted=weighted)
    return model
