"""Settings dictionary and the three ways of equipping a network with a Jacobian output
(jacobianConstruction.py:40-185).  ``equip_*`` return a DinoModel that shares the layers and
weights of ``model`` (as the Keras functional API does) and adds the second output, which the
CUDA library produces with its reverse Jacobian chain instead of TF autodiff.
"""
import numpy as np

from .neuralNetworks import *
from .neuralNetworkUtilities import *
from .trainingUtilities import *
from .dataUtilities import load_data, remap_jacobian_data
from ..model import DinoModel

set_seed(0)     # the reference seeds TF here (jacobianConstruction.py:34-37)


def jacobian_network_settings(problem_settings):
    """Keys and defaults of jacobianConstruction.py:40-99, verbatim."""
    jacobian_settings = {}
    # Jacobian loss
    jacobian_settings['batch_rank'] = 10
    jacobian_settings['target_rank'] = 50
    jacobian_settings['train_full_jacobian'] = False
    # architecture
    jacobian_settings['architecture'] = 'as_dense'
    jacobian_settings['compat_layer'] = True
    jacobian_settings['depth'] = 6
    jacobian_settings['truncation_dimension'] = 50
    jacobian_settings['layer_rank'] = 8
    jacobian_settings['fixed_input_rank'] = 50
    jacobian_settings['fixed_output_rank'] = 50
    jacobian_settings['input_basis'] = 'as'
    jacobian_settings['output_basis'] = 'jjt'
    jacobian_settings['name_prefix'] = 'jacobian_'
    jacobian_settings['breadth_tolerance'] = 1e2
    jacobian_settings['max_breadth'] = 10
    jacobian_settings['input_dim'] = None
    jacobian_settings['output_dim'] = None
    # training
    jacobian_settings['opt_parameters'] = network_training_parameters()
    jacobian_settings['shuffle_every_epoch'] = False
    jacobian_settings['outer_epochs'] = 50
    jacobian_settings['inner_epochs'] = 1
    jacobian_settings['reduced_input_training'] = False
    jacobian_settings['reduced_output_training'] = False
    jacobian_settings['reduced_input_dim'] = None
    jacobian_settings['reduced_output_dim'] = None
    # data
    jacobian_settings['test_data_size'] = 1024
    jacobian_settings['train_data_size'] = 1024
    jacobian_settings['batch_size'] = 256
    jacobian_settings['hess_batch_size'] = 32
    # loading / saving
    jacobian_settings['save_weights'] = True
    jacobian_settings['weights_dir'] = 'trained_weights/'
    jacobian_settings['network_name'] = None
    jacobian_settings['initial_guess_path'] = None
    assert type(problem_settings) is dict, 'problem_settings passed in should have type dict'
    jacobian_settings['problem_settings'] = problem_settings
    return jacobian_settings


def _single_io(model):
    assert len(model.inputs) == 1
    assert len(model.outputs) == 1
    return model.core


def equip_model_with_sketched_jacobian(model, batch_rank, name_prefix=''):
    """Model [m, U, V] -> [q_hat, U^T J_hat V]  with U [B,d_Q,k], V [B,d_M,k], k = batch_rank
    (jacobianConstruction.py:102-125)."""
    return DinoModel(_single_io(model), 'sketched', seed_dim=int(batch_rank), precision=model.precision)


def equip_model_with_full_jacobian(model, name_prefix=''):
    """Model m -> [q_hat, J_hat], J_hat[b,i,j] = d q_hat_i / d m_j  (jacobianConstruction.py:128-153)."""
    return DinoModel(_single_io(model), 'full', precision=model.precision)


def equip_model_with_output_reduced_jacobian(model, reduced_output_basis, name_prefix=''):
    """Model m -> [q_hat, Phi^T J_hat] with Phi = reduced_output_basis [d_Q, r_out]
    (jacobianConstruction.py:155-185)."""
    core = _single_io(model)
    basis = np.asarray(reduced_output_basis)
    assert basis.shape[0] == core.d_out
    return DinoModel(core, 'out_reduced', seed_dim=int(basis.shape[1]), output_basis=basis,
                     precision=model.precision)
