"""Network builders with the reference's names and signatures (neuralNetworks.py:21-172).

They return a :class:`dino_b200.model.DinoModel` instead of a tf.keras Model.  Layer names are
the reference's (they are the keys of the weight pickles), kernels are glorot-uniform and biases
zero like Keras' Dense defaults, and the low-rank residual output layers start at zero
(neuralNetworks.py:112-114).
"""
import math

import numpy as np

from ..model import DinoModel, NetCore

_rng = np.random.default_rng(0)


def set_seed(seed):
    """Reseed the weight initialiser (the drivers' ``tf.random.set_seed(0)``, jacobianConstruction.py:34-37)."""
    global _rng
    _rng = np.random.default_rng(seed)


def _glorot(fan_in, fan_out):
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return _rng.uniform(-lim, lim, size=(fan_in, fan_out)).astype(np.float32)


def _dense(core, units, name, activation=None, use_bias=True, zeros=False, skip_from=-1):
    fan_in = core.n_out[-1] if core.n_out else core.d_in
    kernel = np.zeros((fan_in, units), np.float32) if zeros else _glorot(fan_in, units)
    bias = np.zeros(units, np.float32) if use_bias else None
    return core.add_dense(units, name, activation, use_bias, True, skip_from, kernel, bias)


def _input_stage(input_projector, reduced_input_dim, name_prefix):
    """Input placeholder and the optional frozen projection layer V_r (neuralNetworks.py:27-34)."""
    if input_projector is None:
        assert reduced_input_dim is not None
        return NetCore(reduced_input_dim, name_prefix), reduced_input_dim
    input_dim, reduced_input_dim = input_projector.shape
    core = NetCore(input_dim, name_prefix)
    _dense(core, reduced_input_dim, name_prefix + 'input_proj_layer', use_bias=False)
    return core, reduced_input_dim


def _check_last_layer(last_layer_weights):
    assert type(last_layer_weights) is list
    assert len(last_layer_weights) == 2
    reduced_output_dim, output_dim = last_layer_weights[0].shape
    assert len(last_layer_weights[1].shape) == 1
    assert last_layer_weights[1].shape[0] == output_dim
    return reduced_output_dim, output_dim


def _pin_end_layers(model, input_projector, last_layer_weights, first_layer_trainable, last_layer_trainable,
                    name_prefix):
    if input_projector is not None:
        layer = model.get_layer(name_prefix + 'input_proj_layer')
        layer.trainable = first_layer_trainable
        layer.set_weights([input_projector])
    if last_layer_weights is not None:
        layer = model.get_layer(name_prefix + 'output_layer')
        layer.trainable = last_layer_trainable
        layer.set_weights(last_layer_weights)


def projected_dense(input_projector=None, last_layer_weights=None, hidden_layer_dimensions=[],
                    compat_layer=True, first_layer_trainable=False, last_layer_trainable=False, name_prefix='',
                    reduced_input_dim=None, reduced_output_dim=None):
    """Reduced-basis dense network (neuralNetworks.py:21-73):
    [V_r] -> dense_reduction_layer(softplus) -> inner_layer_i(softplus) -> [output_compat_layer]
    -> [output_layer = (Phi_r^T, mean q)]."""
    core, reduced_input_dim = _input_stage(input_projector, reduced_input_dim, name_prefix)
    if last_layer_weights is not None:
        reduced_output_dim, output_dim = _check_last_layer(last_layer_weights)
    else:
        assert reduced_output_dim is not None
    assert hidden_layer_dimensions[-1] == reduced_output_dim
    _dense(core, reduced_input_dim, name_prefix + 'dense_reduction_layer', 'softplus')
    for i, width in enumerate(hidden_layer_dimensions):
        _dense(core, width, name_prefix + 'inner_layer_' + str(i), 'softplus')
    if compat_layer:
        _dense(core, reduced_output_dim, name_prefix + 'output_compat_layer', use_bias=False)
    if last_layer_weights is None:
        assert compat_layer, 'In this case we need the compat layer'
    else:
        _dense(core, output_dim, name_prefix + 'output_layer')
    regressor = DinoModel(core)
    _pin_end_layers(regressor, input_projector, last_layer_weights, first_layer_trainable, last_layer_trainable,
                    name_prefix)
    return regressor


def generic_dense(input_dim, output_dim, n_hidden_neurons, name_prefix=''):
    """Plain softplus MLP with a linear head (neuralNetworks.py:77-89)."""
    assert type(n_hidden_neurons) is list
    core = NetCore(input_dim, name_prefix)
    for i, width in enumerate(n_hidden_neurons):
        _dense(core, width, name_prefix + 'inner_layer_' + str(i), 'softplus')
    _dense(core, output_dim, name_prefix + 'final_dense')
    return DinoModel(core)


def low_rank_layer(core, rank=8, activation='softplus', name_prefix=None, zeros=True):
    """One low-rank residual branch W2 softplus(W1 z + b1) + b2 appended to ``core``
    (neuralNetworks.py:93-117).  The reference takes a Keras tensor; here the running network is
    the NetCore under construction.  Returns the index of the 'out' layer."""
    width = core.n_out[-1] if core.n_out else core.d_in
    prefix = '' if name_prefix is None else name_prefix
    first = _dense(core, rank, prefix + 'low_rank_residual_in', activation)
    return _dense(core, width, prefix + 'low_rank_residual_out', zeros=zeros, skip_from=first)


def projected_resnet(input_projector, last_layer_weights, ranks=[], compat_layer=True,
                     first_layer_trainable=False, last_layer_trainable=True, name_prefix='',
                     reduced_input_dim=None, reduced_output_dim=None):
    """Reduced-basis ResNet (neuralNetworks.py:122-172), built as the reference intends: its own
    body raises NameError at :142 (undefined hidden_layer_dimensions), rejects the
    reduced_output_dim its caller passes (neuralNetworkUtilities.py:203-205) and has no
    reduced-output variant.  Block i: z <- z + W2 softplus(W1 z + b1) + b2 with W1 [r_out, rank]."""
    core, reduced_input_dim = _input_stage(input_projector, reduced_input_dim, name_prefix)
    if last_layer_weights is not None:
        reduced_output_dim, output_dim = _check_last_layer(last_layer_weights)
    else:
        assert reduced_output_dim is not None
    if not reduced_input_dim == reduced_output_dim:
        _dense(core, reduced_output_dim, name_prefix + 'input_resnet_interface', use_bias=False)
    for i, rank in enumerate(ranks):
        low_rank_layer(core, rank=rank, activation='softplus', name_prefix=name_prefix + str(i))
        core.add_pseudo(name_prefix + 'add' + str(i))
    if compat_layer:
        _dense(core, reduced_output_dim, name_prefix + 'output_compat_layer', use_bias=False)
    if last_layer_weights is not None:
        _dense(core, output_dim, name_prefix + 'output_layer')
    else:
        assert compat_layer, 'In this case we need the compat layer'
    regressor = DinoModel(core)
    _pin_end_layers(regressor, input_projector, last_layer_weights, first_layer_trainable, last_layer_trainable,
                    name_prefix)
    return regressor
