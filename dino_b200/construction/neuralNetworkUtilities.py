"""Reduced-basis selection and architecture dispatch (neuralNetworkUtilities.py:24-233)."""
import numpy as np

from .neuralNetworks import *
from .dataUtilities import get_projectors


def setup_reduced_bases(settings, train_dict, orth_input=True, orth_output=True):
    """Input basis (active subspace 'as' or 'kle') and output basis ('jjt' or 'pod') for the
    rb_* architectures; None otherwise (neuralNetworkUtilities.py:24-75)."""
    if settings['architecture'].lower() not in ['rb_resnet', 'rb_dense']:
        return None
    projectors = get_projectors(settings['data_dir'], fixed_input_rank=settings['fixed_input_rank'],
                                fixed_output_rank=settings['fixed_output_rank'])
    choice = settings['input_basis'].lower()
    if choice == 'as':
        assert 'AS_input' in projectors.keys(), 'Active subspace basis not found'
        input_projector = projectors['AS_input']
    elif choice == 'kle':
        # get_projectors stores the KLE basis under 'KLE' (dataUtilities.py:226); accept both spellings
        key = 'KLE_input' if 'KLE_input' in projectors else 'KLE'
        assert key in projectors.keys(), 'KLE basis not found'
        input_projector = projectors[key]
    else:
        print('Choice of input basis is not supported'.center(80))
        raise ValueError(settings['input_basis'])
    if orth_input:
        input_projector, _ = np.linalg.qr(input_projector)
    choice = settings['output_basis'].lower()
    if choice == 'pod':
        data_dict_pod = {'input_train': train_dict['m_data'], 'output_train': train_dict['q_data']}
        last_layer_weights = build_POD_layer_arrays(data_dict_pod, truncation_dimension=settings['truncation_dimension'],
                                                    breadth_tolerance=settings['breadth_tolerance'],
                                                    max_breadth=settings['max_breadth'])
    elif choice == 'jjt':
        assert 'AS_output' in projectors.keys(), 'Derivative-informed output subspace not found.'
        output_projector = projectors['AS_output']
        if orth_output:
            output_projector, _ = np.linalg.qr(output_projector)
        last_layer_weights = [output_projector.T, np.mean(train_dict['q_data'], axis=0)]   # bias = mean of q
    else:
        print('Choice of output basis is not supported'.center(80))
        raise ValueError(settings['output_basis'])
    return {'input': input_projector, 'output': last_layer_weights[0].T, 'last_layer_bias': last_layer_weights[1]}


def build_POD_layer_arrays(data_dict, truncation_dimension=None, breadth_tolerance=1e2, max_breadth=10):
    """[U_r^T, mean q] from the POD of the centred outputs (neuralNetworkUtilities.py:77-99).  The
    reference reads the singular values before computing them when truncation_dimension is None;
    here the SVD comes first."""
    outputs = data_dict['output_train']
    mean = np.mean(outputs, axis=0)
    full_matrices = truncation_dimension is not None and min(outputs.shape) < truncation_dimension
    if truncation_dimension is not None:
        assert type(truncation_dimension) is int
    U, s, _ = np.linalg.svd((outputs - mean).T, full_matrices=full_matrices)
    if truncation_dimension is None:
        orders_reduction = s[0] / (s + 1e-12)
        abs_tol_idx = np.where(orders_reduction > breadth_tolerance)[0][0]
        truncation_dimension = min(abs_tol_idx, max_breadth)
    return [U[:, :truncation_dimension].T, mean]


def choose_network(settings, projector_dict=None, reduced_input_training=False, reduced_output_training=False):
    """Build the network named by settings['architecture']: 'rb_resnet', 'rb_dense' or
    'generic_dense' (neuralNetworkUtilities.py:102-195)."""
    depth = settings['depth']
    architecture = settings['architecture']
    prefix = settings['name_prefix']
    if architecture.lower() in ['rb_resnet', 'rb_dense']:
        assert projector_dict is not None
        print(80 * '#')
        kind = 'ResNet' if architecture.lower() == 'rb_resnet' else 'Dense'
        print(('Loading ' + settings['input_basis'].upper() + '-' + settings['output_basis'].upper() + ' ' + kind).center(80))
        if reduced_input_training:
            input_projector = None
            assert settings['reduced_input_dim'] is not None
        else:
            input_projector = projector_dict['input']
        if reduced_output_training:
            last_layer_weights = None
            assert settings['reduced_output_dim'] is not None
        else:
            last_layer_weights = [projector_dict['output'].T, projector_dict['last_layer_bias']]
        common = dict(name_prefix=prefix, reduced_input_dim=settings['reduced_input_dim'],
                      reduced_output_dim=settings['reduced_output_dim'], compat_layer=settings['compat_layer'])
        if architecture.lower() == 'rb_resnet':
            regressor, _ = construct_projected_resnet(input_projector, last_layer_weights,
                                                      depth * [settings['layer_rank']], **common)
        else:
            regressor, _ = construct_projected_dense(input_projector, last_layer_weights, depth, **common)
    elif architecture == 'generic_dense':
        print(80 * '#')
        print('Loading generic dense'.center(80))
        assert settings['input_dim'] is not None
        assert settings['output_dim'] is not None
        regressor = generic_dense(settings['input_dim'], settings['output_dim'], depth * [settings['truncation_dimension']])
    else:
        raise Exception('Architecture: ', architecture, ' not supported!')
    return regressor


def construct_projected_resnet(input_projector, last_layer_weights, ranks, name_prefix='',
                               reduced_input_dim=None, reduced_output_dim=None, compat_layer=True):
    """Reduced-basis ResNet (neuralNetworkUtilities.py:198-207)."""
    pod_resnet = projected_resnet(input_projector=input_projector, last_layer_weights=last_layer_weights, ranks=ranks,
                                  name_prefix=name_prefix, reduced_input_dim=reduced_input_dim,
                                  reduced_output_dim=reduced_output_dim, compat_layer=compat_layer)
    return pod_resnet, last_layer_weights


def construct_projected_dense(input_projector, last_layer_weights, depth, name_prefix='', reduced_input_dim=None,
                              reduced_output_dim=None, truncation_dimension=None, compat_layer=True):
    """Reduced-basis dense network whose hidden width is inferred the reference's way: the reduced
    output dimension if the last layer is given, else the reduced input dimension, else the
    reduced output dimension (neuralNetworkUtilities.py:209-233)."""
    if truncation_dimension is None:
        if last_layer_weights is not None:
            truncation_dimension = last_layer_weights[0].shape[0]
        elif reduced_input_dim is not None:
            print('Assuming that truncation_dimension is reduced input dimension')
            truncation_dimension = reduced_input_dim
        elif reduced_output_dim is not None:
            print('Assuming that truncation_dimension is reduced output dimension')
            truncation_dimension = reduced_output_dim
        else:
            print('truncation_dimension must be specified, or inferrable in some way.')
            raise ValueError('truncation_dimension')
    pod_dense_network = projected_dense(input_projector=input_projector, last_layer_weights=last_layer_weights,
                                        hidden_layer_dimensions=depth * [truncation_dimension], name_prefix=name_prefix,
                                        reduced_input_dim=reduced_input_dim, reduced_output_dim=reduced_output_dim,
                                        compat_layer=compat_layer)
    return pod_dense_network, last_layer_weights
