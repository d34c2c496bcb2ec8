"""End-to-end DINO construction: data -> reduced bases -> pruning -> network -> H1 training ->
re-stitching and weight pickle (jacobianTrainingDriver.py:19-396).  Same five stages, same
function names, settings keys and side effects on the data dictionaries as the reference.
"""
import os
import pickle

import numpy as np

from .neuralNetworks import *
from .neuralNetworkUtilities import *
from .trainingUtilities import *
from .dataUtilities import load_data, remap_jacobian_data, flatten_data, train_test_split
from .jacobianConstruction import (equip_model_with_full_jacobian, equip_model_with_output_reduced_jacobian,
                                   equip_model_with_sketched_jacobian)
from ..model import Adam


def _default_data_dir(settings):
    ps = settings['problem_settings']
    return '../data/' + ps['formulation'] + '_n_obs_' + str(ps['ntargets']) + '_g' + str(ps['gamma']) + \
        '_d' + str(ps['delta']) + '_nx' + str(ps['nx']) + '/'


def get_the_data(settings, remapped_data=None, unflattened_data=None, verbose=True):
    """Training / testing dictionaries (jacobianTrainingDriver.py:19-90).  Returns
    {'train_dict', 'test_dict', 'settings', 'unflattened_train_dict'} and records the full
    input/output dimensions in the settings."""
    unflattened_train_dict = None
    if settings['data_dir'] is None:
        settings['data_dir'] = _default_data_dir(settings)
    data_dir = settings['data_dir']
    assert os.path.isdir(data_dir), 'Directory does not exist' + data_dir
    if remapped_data is not None:
        assert settings['train_full_jacobian']
        train_dict = remapped_data[0]
        test_dict = None if len(remapped_data) == 1 else remapped_data[1]
    else:
        if unflattened_data is None:
            n_data = settings['train_data_size'] + settings['test_data_size']
            all_data = load_data(data_dir, rescale=False, n_data=n_data, derivatives=True)
            unflattened_train_dict, unflattened_test_dict = train_test_split(all_data, n_train=settings['train_data_size'])
        else:
            unflattened_train_dict = unflattened_data[0]
            unflattened_test_dict = None if len(unflattened_data) == 1 else unflattened_data[1]
        if settings['train_full_jacobian']:
            # settings['device_data'] (extension, default False): form the dense targets with device contractions
            # (float32) instead of the reference's host float64 loop
            if settings.get('device_data', False):
                from .deviceData import remap_jacobian_data_device as shape_targets
            else:
                shape_targets = remap_jacobian_data
        else:
            def shape_targets(d):
                return flatten_data(d, target_rank=settings['target_rank'], batch_rank=settings['batch_rank'])
        train_dict = shape_targets(unflattened_train_dict)
        test_dict = None if unflattened_test_dict is None else shape_targets(unflattened_test_dict)
    assert train_dict['m_data'].shape[0] != 0
    assert test_dict['m_data'].shape[0] != 0          # as in the reference: a test set is required here
    assert len(train_dict['m_data'].shape) == 2
    assert len(train_dict['q_data'].shape) == 2
    settings['input_dim'] = train_dict['m_data'].shape[1]
    settings['output_dim'] = (test_dict if test_dict is not None else train_dict)['q_data'].shape[-1]
    return {'train_dict': train_dict, 'test_dict': test_dict, 'settings': settings,
            'unflattened_train_dict': unflattened_train_dict}


def prune_the_data(projector_dict, train_dict, test_dict, input_pruning=True, output_pruning=True, verbose=False):
    """Restrict the data to the reduced bases (jacobianTrainingDriver.py:93-185):
    m <- m V, J <- J V; q <- (q - b) Phi, J <- Phi^T J.  The unreduced arrays are kept under
    'm_full' / 'q_full' / 'J_full' for the re-stitched evaluation."""
    assert input_pruning or output_pruning
    dicts = (train_dict, test_dict)
    if 'J_data' in train_dict.keys():
        J = [d['J_data'] for d in dicts]
        J_output_already_pruned = False
    elif 'PhiTJ_data' in train_dict.keys():
        J = [d['PhiTJ_data'] for d in dicts]
        J_output_already_pruned = True
        print('Assuming that Jacobian already pruned wrt outputs!')
    else:
        raise KeyError('J_data')
    for d, Jd in zip(dicts, J):
        d['J_full'] = Jd.copy()
    if input_pruning:
        V = projector_dict['input']
        assert V.shape[0] == train_dict['m_data'].shape[1]
        for i, d in enumerate(dicts):
            d['m_full'] = d['m_data'].copy()
            d['m_data'] = d['m_data'] @ V
            J[i] = J[i] @ V
    if output_pruning:
        Phi, bias = projector_dict['output'], projector_dict['last_layer_bias']
        assert Phi.shape[0] == train_dict['q_data'].shape[1]
        for i, d in enumerate(dicts):
            d['q_full'] = d['q_data'].copy()
            d['q_data'] = (d['q_data'] - bias) @ Phi
            if J_output_already_pruned:
                assert J[i].shape[1] == Phi.shape[1]
            else:
                J[i] = np.matmul(Phi.T, J[i])
    for d, Jd in zip(dicts, J):
        d['J_data'] = Jd
    if verbose:
        for k in ('m_data', 'q_data', 'J_data'):
            print(k, train_dict[k].shape, test_dict[k].shape)
    return train_dict, test_dict


def setup_the_dino(settings, train_dict, projector_dict=None, reduced_input_training=False,
                   reduced_output_training=False, no_jacobian=False, reduced_output_Jacobian=False):
    """Build the network, optionally initialise it from a weight pickle, and equip it with the
    Jacobian output the settings ask for (jacobianTrainingDriver.py:188-224)."""
    regressor = choose_network(settings, projector_dict, reduced_input_training=reduced_input_training,
                               reduced_output_training=reduced_output_training)
    if settings['initial_guess_path'] is not None:
        path = settings['initial_guess_path']
        assert os.path.isfile(path), 'Trained weights may not exist as specified: ' + str(path)
        with open(path, 'rb') as f:
            regressor_weights = pickle.load(f)
        for layer in regressor.layers:
            layer.set_weights(regressor_weights[layer.name])
        if settings['opt_parameters']['train_hessianlearn']:
            settings['opt_parameters']['layer_weights'] = regressor_weights
    if not no_jacobian:
        if reduced_output_Jacobian:
            regressor = equip_model_with_output_reduced_jacobian(regressor, projector_dict['output'],
                                                                 name_prefix=settings['name_prefix'])
        elif settings['train_full_jacobian']:
            print('Equipping Jacobian')
            settings['opt_parameters']['train_full_jacobian'] = settings['train_full_jacobian']
            regressor = equip_model_with_full_jacobian(regressor, name_prefix=settings['name_prefix'])
        else:
            regressor = equip_model_with_sketched_jacobian(regressor, settings['batch_rank'],
                                                           name_prefix=settings['name_prefix'])
    return regressor


def train_dino(settings, regressor, train_dict, test_dict, unflattened_train_dict=None, verbose=True):
    """Run the H1 training; with shuffle_every_epoch the sketch columns are re-drawn before every
    outer epoch (jacobianTrainingDriver.py:227-246)."""
    print('Commencing training'.center(80))
    if settings['shuffle_every_epoch']:
        assert unflattened_train_dict is not None
        settings['opt_parameters']['keras_epochs'] = settings['inner_epochs']
        for epoch in range(settings['outer_epochs']):
            if verbose:
                print(('Running inner iteration ' + str(epoch)).center(80))
            if settings.get('device_data', False):
                # the factors go to the device once; every outer epoch re-draws and regroups the sketch columns there
                if epoch == 0:
                    from .deviceData import upload_factors, flatten_data_device
                    device_factors = upload_factors(unflattened_train_dict)
                train_dict = flatten_data_device(device_factors, target_rank=settings['target_rank'],
                                                 batch_rank=settings['batch_rank'], order_random=True, burn_in=epoch)
            else:
                train_dict = flatten_data(unflattened_train_dict, target_rank=settings['target_rank'],
                                          batch_rank=settings['batch_rank'], order_random=True, burn_in=epoch)
            regressor = train_h1_network(regressor, train_dict, test_dict, opt_parameters=settings['opt_parameters'],
                                         verbose=True)
    else:
        regressor = train_h1_network(regressor, train_dict, test_dict, opt_parameters=settings['opt_parameters'])
    return regressor


def restitch_and_postprocess(reduced_regressor, settings, train_dict, test_dict, projector_dict, l2_only=False,
                             reduced_output_Jacobian=False, verbose=False, logger={}):
    """Rebuild the full network around the reduced one (weights copied by layer name), evaluate it on
    the unreduced data and pickle {layer.name: weights} (jacobianTrainingDriver.py:248-342)."""
    for _ in range(5):
        print(80 * '#')
    print('Re-stitching post processing')
    regressor = setup_the_dino(settings, train_dict, projector_dict, reduced_input_training=False,
                               reduced_output_training=False, no_jacobian=l2_only,
                               reduced_output_Jacobian=reduced_output_Jacobian)
    big_network_layers = [layer.name for layer in regressor.layers]
    for layer in reduced_regressor.layers:
        if layer.name in big_network_layers:
            regressor.get_layer(layer.name).set_weights(layer.get_weights())
    opt_parameters = settings['opt_parameters']
    optimizer = Adam(learning_rate=opt_parameters['keras_alpha'])
    if l2_only:
        assert len(regressor.outputs) == 1
        regressor.compile(optimizer=optimizer, loss=normalized_mse, loss_weights=[1.0], metrics=[l2_accuracy])
    else:
        regressor.compile(optimizer=optimizer, loss=[normalized_mse, normalized_mse_matrix],
                          loss_weights=opt_parameters['loss_weights'], metrics=[l2_accuracy, f_accuracy_matrix])
    if not ('q_full' in train_dict.keys()):
        train_dict['q_full'] = train_dict['q_data']
        test_dict['q_full'] = test_dict['q_data']
    input_train, input_test = [train_dict['m_full']], [test_dict['m_full']]
    if l2_only:
        output_train, output_test = [train_dict['q_full']], [test_dict['q_full']]
    else:
        output_train = [train_dict['q_full'], train_dict['J_full']]
        output_test = [test_dict['q_full'], test_dict['J_full']]
    eval_train = regressor.evaluate(input_train, output_train, verbose=2)
    eval_test = regressor.evaluate(input_test, output_test, verbose=2)
    if l2_only:
        print('After training: l2 accuracy = ', eval_train[1])
        print('After training: test l2 accuracy = ', eval_test[1])
        logger['l2_train'], logger['l2_test'] = eval_train[1], eval_test[1]
    else:
        print('After training: l2, h1 training accuracies = ', eval_train[3], eval_train[6])
        if verbose:
            print('eval_test_dict = ', dict(zip(regressor.metrics_names, eval_test)))
        print('After training: l2, h1 testing accuracies =  ', eval_test[3], eval_test[6])
        logger['l2_train'], logger['h1_train'] = eval_train[3], eval_train[6]
        logger['l2_test'], logger['h1_test'] = eval_test[3], eval_test[6]
    if settings['save_weights']:
        jacobian_weights = {layer.name: layer.get_weights() for layer in regressor.layers}
        os.makedirs(settings['weights_dir'], exist_ok=True)
        if settings['network_name'] is None:
            network_name = settings['name_prefix'] + str(settings['architecture']) + '_depth' + str(settings['depth']) + \
                '_batch_rank_' + str(settings['batch_rank'])
        else:
            network_name = settings['network_name']
        with open(settings['weights_dir'] + network_name + '.pkl', 'wb+') as f_jacobian:
            pickle.dump(jacobian_weights, f_jacobian, pickle.HIGHEST_PROTOCOL)
    return regressor


def jacobian_training_driver(settings, remapped_data=None, unflattened_data=None, verbose=True):
    """The complete DINO construction (jacobianTrainingDriver.py:346-396)."""
    for loss_weight in settings['opt_parameters']['loss_weights']:
        assert loss_weight >= 0
    data_dict = get_the_data(settings, remapped_data=remapped_data, unflattened_data=unflattened_data, verbose=verbose)
    train_dict, test_dict = data_dict['train_dict'], data_dict['test_dict']
    unflattened_train_dict = data_dict['unflattened_train_dict']
    projector_dict = setup_reduced_bases(settings, train_dict)
    reduced = settings['reduced_input_training'] or settings['reduced_output_training']
    if reduced:
        assert settings['train_full_jacobian']
        assert len(projector_dict['input'].shape) == 2          # (dM, rM), assumed orthonormal
        if settings['reduced_input_training']:
            settings['reduced_input_dim'] = projector_dict['input'].shape[1]
            print('reduced_input_dim = ', settings['reduced_input_dim'])
        if settings['reduced_output_training']:
            settings['reduced_output_dim'] = projector_dict['output'].shape[1]
        if settings.get('device_data', False):
            from .deviceData import prune_the_data_device as prune
        else:
            prune = prune_the_data
        train_dict, test_dict = prune(projector_dict, train_dict, test_dict,
                                      input_pruning=settings['reduced_input_training'],
                                      output_pruning=settings['reduced_output_training'])
    regressor = setup_the_dino(settings, train_dict, projector_dict,
                               reduced_input_training=settings['reduced_input_training'],
                               reduced_output_training=settings['reduced_output_training'])
    regressor = train_dino(settings, regressor, train_dict, test_dict, unflattened_train_dict)
    if reduced:
        regressor = restitch_and_postprocess(regressor, settings, train_dict, test_dict, projector_dict)
    return regressor
