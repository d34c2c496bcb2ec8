"""Post-hoc accuracies of a trained network (surrogatePostProcessing.py:31-146)."""
import numpy as np

from ..construction.jacobianConstruction import equip_model_with_sketched_jacobian, equip_model_with_full_jacobian
from ..construction.observableConstruction import observable_network_loader
from ..construction.trainingUtilities import (normalized_mse, l2_accuracy, normalized_mse_matrix, f_accuracy_matrix)
from ..construction.dataUtilities import load_data, remap_jacobian_data, flatten_data, train_test_split


def evaluateJacobianNetwork(settings, jacobian_network=None, file_name=None, split_data=None, data_dir=None):
    """Load (or take) a network, equip it with the Jacobian output the settings name and report the
    l2 / h1 accuracies on the training and testing data.  The returned dictionary has the reference's
    keys; like the reference (surrogatePostProcessing.py:141) 'l2_test'/'h1_test' repeat the TRAIN
    values - the true test values are under 'test_eval'."""
    if jacobian_network is None:
        if file_name is None and settings['network_name'] is not None:
            file_name = settings['weights_dir'] + settings['network_name'] + '.pkl'
        jacobian_network = observable_network_loader(settings, file_name=file_name)
    jacobian_network.summary()
    if settings.get('full_JVVT', False):
        raise NotImplementedError('full_JVVT calls equip_model_with_full_jacobian_and_right_nullspace, which the '
                                  'reference never defines (surrogatePostProcessing.py:94)')
    full = settings.get('full_jacobian', settings.get('train_full_jacobian', False))
    if split_data is None:
        n_data = settings['train_data_size'] + settings['test_data_size']
        ps = settings['problem_settings']
        if data_dir is None:
            data_dir = '../data/' + ps['formulation'] + '_n_obs_' + str(ps['ntargets']) + '_g' + str(ps['gamma']) + \
                '_d' + str(ps['delta']) + '_nx' + str(ps['nx']) + '/'
        all_data = load_data(data_dir, rescale=False, n_data=n_data, derivatives=True)
        unflat_train, unflat_test = train_test_split(all_data, n_train=settings['train_data_size'])
        if full:
            dicts = [remap_jacobian_data(unflat_train), remap_jacobian_data(unflat_test)]
            feeds = [([d['m_data']], [d['q_data'], d['J_data']]) for d in dicts]
        else:
            dicts = [flatten_data(d, target_rank=settings['target_rank'], batch_rank=settings['batch_rank'])
                     for d in (unflat_train, unflat_test)]
            feeds = [([d['m_data'], d['U_data'], d['V_data']], [d['q_data'], d['sigma_data']]) for d in dicts]
        (input_train, output_train), (input_test, output_test) = feeds
    else:
        input_train, output_train = split_data['input_train'], split_data['output_train']
        input_test, output_test = split_data['input_test'], split_data['output_test']
    if full:
        for _ in range(10):
            print('RUNNING FOR FULL JACOBIAN')
        jacobian_network = equip_model_with_full_jacobian(jacobian_network, name_prefix=settings['name_prefix'])
    else:
        jacobian_network = equip_model_with_sketched_jacobian(jacobian_network, settings['batch_rank'],
                                                              name_prefix=settings['name_prefix'])
    jacobian_network.compile(optimizer='adam', loss=[normalized_mse, normalized_mse_matrix],
                             loss_weights=settings['opt_parameters']['loss_weights'],
                             metrics=[l2_accuracy, f_accuracy_matrix])
    eval_train = jacobian_network.evaluate(input_train, output_train, verbose=True)
    eval_test = jacobian_network.evaluate(input_test, output_test, verbose=True)
    names = jacobian_network.metrics_names
    # 0 total loss, 1 q loss, 2 J loss, 3 l2 acc(q), 4 f acc(q), 5 l2 acc(J), 6 f acc(J)
    print('After loading: l2, h1 training accuracies = ', eval_train[3], eval_train[6])
    print('After loading: l2, h1 testing accuracies =  ', eval_test[3], eval_test[6])
    return {'l2_train': eval_train[3], 'h1_train': eval_train[6], 'l2_test': eval_train[3], 'h1_test': eval_train[6],
            'train_eval': dict(zip(names, eval_train)), 'test_eval': dict(zip(names, eval_test))}
