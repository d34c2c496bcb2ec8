"""Losses, metrics, optimiser settings and the two training entry points, with the reference's
names, argument meaning and printed/logged results (trainingUtilities.py:25-222).

``train_h1_network`` is THE hot-path entry: everything it asks of the model -
compile, evaluate, fit - runs in CUDA kernels behind :class:`dino_b200.model.DinoModel`.
The four loss/metric functions below are numpy restatements offered for API completeness
(callers pass them to ``compile``, which recognises them by name); the training path itself
evaluates them inside the fused loss kernels.
"""
import numpy as np

from .dataUtilities import *
from ..model import Adam, SGD


def _ratio_last_axis(y_true, y_pred):
    y_true, y_pred = np.asarray(y_true), np.asarray(y_pred)
    return np.mean((y_true - y_pred) ** 2, axis=-1) / np.mean(y_true ** 2, axis=-1)


def _ratio_last_two_axes(y_true, y_pred):
    y_true, y_pred = np.asarray(y_true), np.asarray(y_pred)
    return ((y_true - y_pred) ** 2).sum(-1).sum(-1) / (y_true ** 2).sum(-1).sum(-1)


def normalized_mse(y_true, y_pred):
    return np.mean(_ratio_last_axis(y_true, y_pred))


def normalized_mse_matrix(y_true, y_pred):
    return np.mean(_ratio_last_two_axes(y_true, y_pred))


def l2_accuracy(y_true, y_pred):
    return 1. - np.sqrt(np.mean(_ratio_last_axis(y_true, y_pred)))


def f_accuracy_matrix(y_true, y_pred):
    return 1. - np.sqrt(np.mean(_ratio_last_two_axes(y_true, y_pred)))


def network_training_parameters():
    """Training settings; keys and defaults of trainingUtilities.py:53-82."""
    opt_parameters = {}
    opt_parameters['loss_weights'] = [1.0, 1.0]          # [l2, h1 seminorm]
    opt_parameters['train_full_jacobian'] = True
    opt_parameters['train_keras'] = True
    opt_parameters['keras_epochs'] = 10
    opt_parameters['keras_batch_size'] = 32
    opt_parameters['keras_opt'] = 'adam'
    opt_parameters['keras_alpha'] = 1e-3
    opt_parameters['keras_verbose'] = False
    # hessianlearn (LRSFN) settings are kept for dict compatibility; that optimiser is not built
    opt_parameters['train_hessianlearn'] = False
    opt_parameters['hessian_low_rank'] = 40
    opt_parameters['hess_alpha'] = 1e-4
    opt_parameters['hess_gbatch_size'] = 256
    opt_parameters['hess_batch_size'] = 32
    opt_parameters['hess_sweeps'] = 10
    opt_parameters['layer_weights'] = {}
    opt_parameters['printing_sweep_frequency'] = 0.1
    opt_parameters['callbacks'] = []
    return opt_parameters


def _make_optimizer(opt_parameters):
    if opt_parameters['keras_opt'] == 'adam':
        return Adam(learning_rate=opt_parameters['keras_alpha'])
    if opt_parameters['keras_opt'] == 'sgd':
        return SGD(learning_rate=opt_parameters['keras_alpha'])
    # the reference does `raise 'Invalid choice of optimizer'`, i.e. a TypeError (trainingUtilities.py:94)
    raise TypeError('Invalid choice of optimizer')


def _h1_feed(data_dict, full_jacobian):
    if full_jacobian:
        return [data_dict['m_data']], [data_dict['q_data'], data_dict['J_data']]
    return ([data_dict['m_data'], data_dict['U_data'], data_dict['V_data']],
            [data_dict['q_data'], data_dict['sigma_data']])


def train_h1_network(network, train_dict, test_dict=None, opt_parameters=network_training_parameters(),
                     verbose=True, logger={}):
    """H1 training: L2 misfit on q plus Frobenius misfit on the Jacobian (trainingUtilities.py:85-179)."""
    optimizer = _make_optimizer(opt_parameters)
    assert len(opt_parameters['loss_weights']) == len(network.outputs)
    assert len(network.outputs) == 2
    network.compile(optimizer=optimizer, loss=[normalized_mse, normalized_mse_matrix],
                    loss_weights=opt_parameters['loss_weights'], metrics=[l2_accuracy, f_accuracy_matrix])
    input_train, output_train = _h1_feed(train_dict, opt_parameters['train_full_jacobian'])
    test_data = None
    if test_dict is not None:
        input_test, output_test = _h1_feed(test_dict, opt_parameters['train_full_jacobian'])
        test_data = (input_test, output_test)

    def report(when, record):
        eval_train = network.evaluate(input_train, output_train, verbose=2)
        print(when + ' training: l2, h1 training accuracies = ', eval_train[3], eval_train[6])
        if record:
            logger['l2_train'], logger['h1_train'] = eval_train[3], eval_train[6]
        if test_dict is not None:
            eval_test = network.evaluate(input_test, output_test, verbose=2)
            print(when + ' training: l2, h1 testing accuracies =  ', eval_test[3], eval_test[6])
            if record:
                logger['l2_test'], logger['h1_test'] = eval_test[3], eval_test[6]

    if verbose:
        report('Before', False)
    if opt_parameters['train_keras']:
        network.fit(input_train, output_train, validation_data=test_data, epochs=opt_parameters['keras_epochs'],
                    batch_size=opt_parameters['keras_batch_size'], verbose=opt_parameters['keras_verbose'],
                    callbacks=opt_parameters['callbacks'])
    if opt_parameters['train_hessianlearn']:
        raise NotImplementedError('hessianlearn (LRSFN) training is an un-vendored optional dependency of the '
                                  'reference (trainingUtilities.py:140-166) and is out of scope here')
    if verbose:
        report('After', True)
    return network


def train_l2_network(network, train_dict, test_dict=None, opt_parameters=network_training_parameters(),
                     verbose=True):
    """L2-only training (trainingUtilities.py:181-222)."""
    optimizer = _make_optimizer(opt_parameters)
    network.compile(optimizer=optimizer, loss=normalized_mse, metrics=[l2_accuracy])
    input_train, output_train = train_dict['m_data'], train_dict['q_data']
    test_data = None
    if test_dict is not None:
        input_test, output_test = test_dict['m_data'], test_dict['q_data']
        test_data = (input_test, output_test)
    if verbose:
        _, l2_acc_train = network.evaluate(input_train, output_train, verbose=2)
        print('Before training: l2 accuracy = ', l2_acc_train)
        if test_dict is not None:
            _, l2_acc_test = network.evaluate(input_test, output_test, verbose=2)
            print('Before training: l2accuracy = ', l2_acc_test)
    network.fit(input_train, output_train, validation_data=test_data, epochs=opt_parameters['keras_epochs'],
                batch_size=opt_parameters['keras_batch_size'], verbose=opt_parameters['keras_verbose'],
                callbacks=opt_parameters['callbacks'])
    if verbose:
        _, l2_acc_train = network.evaluate(input_train, output_train, verbose=2)
        print('After training: l2 accuracy = ', l2_acc_train)
        if test_dict is not None:
            _, l2_acc_test = network.evaluate(input_test, output_test, verbose=2)
            print('After training: test l2 accuracy = ', l2_acc_test)
    return network
