"""L2-only ("observable") networks: settings, driver, and the weight-pickle loader that also
serves DINO networks (observableConstruction.py:38-165)."""
import os
import pickle

import numpy as np

from .neuralNetworks import *
from .neuralNetworkUtilities import *
from .trainingUtilities import *
from .dataUtilities import load_data

set_seed(0)


def observable_network_settings(problem_settings):
    """Keys and defaults of observableConstruction.py:38-78."""
    settings = {}
    settings['architecture'] = 'rb_dense'
    settings['compat_layer'] = True
    settings['depth'] = 6
    settings['truncation_dimension'] = 50
    settings['layer_rank'] = 8
    settings['fixed_input_rank'] = 50
    settings['fixed_output_rank'] = 50
    settings['input_basis'] = 'as'
    settings['output_basis'] = 'jjt'
    settings['name_prefix'] = 'observable_'
    settings['breadth_tolerance'] = 1e2
    settings['max_breadth'] = 10
    settings['weights_dir'] = 'trained_weights/'
    settings['opt_parameters'] = network_training_parameters()
    settings['epochs'] = 50
    settings['test_data_size'] = 1024
    settings['train_data_size'] = 15 * 1024
    settings['batch_size'] = 256
    settings['hess_batch_size'] = 32
    assert type(problem_settings) is dict, 'problem_settings passed in should have type dict'
    settings['problem_settings'] = problem_settings
    return settings


def observable_training_driver(settings, verbose=True):
    """Plain L2 training of m -> q (observableConstruction.py:81-116)."""
    n_data = settings['train_data_size'] + settings['test_data_size']
    if settings['data_dir'] is None:
        ps = settings['problem_settings']
        settings['data_dir'] = '../data/' + ps['formulation'] + '_n_obs_' + str(ps['ntargets']) + '_g' + \
            str(ps['gamma']) + '_d' + str(ps['delta']) + '_nx' + str(ps['nx']) + '/'
    data_dir = settings['data_dir']
    assert os.path.isdir(data_dir), 'Directory does not exist' + data_dir
    for loss_weight in settings['opt_parameters']['loss_weights']:
        assert loss_weight >= 0
    all_data = load_data(data_dir, rescale=False, n_data=n_data, derivatives=False)
    settings.setdefault('input_dim', all_data['m_data'].shape[-1])
    settings.setdefault('output_dim', all_data['q_data'].shape[-1])
    train_dict, test_dict = train_test_split(all_data, n_train=settings['train_data_size'])
    if settings['architecture'].lower() in ['rb_resnet', 'rb_dense']:
        projector_dict = setup_reduced_bases(settings, train_dict)
    else:
        projector_dict = None
    regressor = choose_network(settings, projector_dict)
    print('Commencing training'.center(80))
    return train_l2_network(regressor, train_dict, test_dict, opt_parameters=settings['opt_parameters'])


def observable_network_loader(settings, file_name=None):
    """Rebuild a network from a {layer name: [arrays]} pickle; the projectors and the presence of
    the compat layer are inferred from the keys (observableConstruction.py:119-164)."""
    if file_name is None:
        file_name = settings['weights_dir'] + settings['name_prefix'] + str(settings['architecture']) + \
            '_depth' + str(settings['depth']) + '.pkl'
    assert os.path.isfile(file_name), 'Trained weights may not exist as specified: ' + str(file_name)
    with open(file_name, 'rb') as f:
        observable_weights = pickle.load(f)
    prefix = settings['name_prefix']
    try:
        projector_dict = {'input': observable_weights[prefix + 'input_proj_layer'][0],
                          'output': observable_weights[prefix + 'output_layer'][0].T,
                          'last_layer_bias': observable_weights[prefix + 'output_layer'][1]}
    except KeyError:
        projector_dict = {}
    settings['compat_layer'] = any('compat_layer' in name for name in observable_weights)
    observable_network = choose_network(settings, projector_dict)
    for layer in observable_network.layers:
        layer.set_weights(observable_weights[layer.name])
    return observable_network
