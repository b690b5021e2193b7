"""Shared test scaffolding: the cfg sections used for the golden fixtures and builders for the
product's plugin classes."""
import os

import numpy as np
import torch

from conftest import GOLDEN

LSTM_OPTS = dict(
    arch_name='LSTM_layers', arch_proto='proto/LSTM.proto', arch_library='neural_networks', arch_class='LSTM',
    arch_pretrain_file='none', arch_freeze='False', arch_seq_model='True', lstm_lay='24,24', lstm_drop='0.2,0.2',
    lstm_use_laynorm_inp='False', lstm_use_batchnorm_inp='False', lstm_use_laynorm='False,False',
    lstm_use_batchnorm='True,True', lstm_bidir='False', lstm_act='tanh,tanh', if_hsigmoid='True',
    lstm_orthinit='True', arch_lr='0.0016', arch_halving_factor='0.5', arch_improvement_threshold='0.001',
    arch_opt='rmsprop', opt_momentum='0.0', opt_alpha='0.95', opt_eps='1e-8', opt_centered='False',
    opt_weight_decay='0.0', out_folder='', lstm_hcgs='True', guided_hcgs='False', apply_guided_hcgs='False',
    hcgsx_block='8,4', hcgsx_sparse='50,50', hcgsh_block='8,4', hcgsh_sparse='50,50', lstm_quant='False',
    param_quant='8,8', lstm_quant_inp='False', inp_quant='16', lstm_prune='False', lstm_prune_perc='70',
    skip_regularization='True')

MLP_OPTS = dict(
    arch_name='MLP_layers', arch_proto='proto/MLP.proto', arch_library='neural_networks', arch_class='MLP',
    arch_pretrain_file='none', arch_freeze='False', arch_seq_model='False', dnn_lay='10', dnn_drop='0.0',
    dnn_use_laynorm_inp='False', dnn_use_batchnorm_inp='False', dnn_use_batchnorm='False', dnn_use_laynorm='False',
    dnn_act='softmax', arch_lr='0.0016', arch_halving_factor='0.5', arch_improvement_threshold='0.001',
    arch_opt='rmsprop', opt_momentum='0.0', opt_alpha='0.95', opt_eps='1e-8', opt_centered='False',
    opt_weight_decay='0.0', out_folder='', mlp_hcgs='True', guided_hcgs='False', apply_guided_hcgs='False',
    hcgs_block='8,4', hcgs_sparse='50,25', mlp_quant='False', param_quant='8', mlp_quant_inp='False',
    inp_quant='16', mlp_prune='False', mlp_prune_perc='70', skip_regularization='True')

MODEL_2HEAD = ['out_dnn1=compute(LSTM_layers,fbank)', 'out_dnn2=compute(MLP_layers,out_dnn1)',
               'out_dnn3=compute(MLP_layers2,out_dnn1)', 'loss_mono=cost_nll(out_dnn3,lab_mono)',
               'loss_mono_w=mult_constant(loss_mono,1.0)', 'loss_cd=cost_nll(out_dnn2,lab_cd)',
               'loss_final=sum(loss_cd,loss_mono_w)', 'err_final=cost_err(out_dnn2,lab_cd)']

NETS = {'lstm': 'LSTM_layers', 'cd': 'MLP_layers', 'mono': 'MLP_layers2'}


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False)


def section_opts(to_do, lstm_over=None, mlp_over=None, out_folder='', use_cuda='False', mono_over=None):
    a1 = dict(LSTM_OPTS, out_folder=out_folder, to_do=to_do, use_cuda=use_cuda)
    a1.update(lstm_over or {})
    a2 = dict(MLP_OPTS, out_folder=out_folder, to_do=to_do, use_cuda=use_cuda)
    a2.update(mlp_over or {})
    a3 = dict(a2, arch_name='MLP_layers2', dnn_lay='5', mlp_hcgs='False', arch_lr='0.0004')
    a3.update(mono_over or {})
    return a1, a2, a3


def build_product_nets(to_do, D=12, seed=2234, lstm_over=None, mlp_over=None, out_folder='', device='cpu',
                       mono_over=None):
    """Construct LSTM + two MLP heads in the cfg's ``model=`` order with the run_nn seeding."""
    import neural_networks as nn_b200
    torch.manual_seed(seed)
    np.random.seed(seed)
    a1, a2, a3 = section_opts(to_do, lstm_over, mlp_over, out_folder, use_cuda=str(device != 'cpu'),
                              mono_over=mono_over)
    lstm = nn_b200.LSTM(a1, D)
    cd = nn_b200.MLP(a2, lstm.out_dim)
    mono = nn_b200.MLP(a3, lstm.out_dim)
    nets = {'LSTM_layers': lstm, 'MLP_layers': cd, 'MLP_layers2': mono}
    for n in nets.values():
        n.to(device)
        n.train() if to_do == 'train' else n.eval()
    return nets, (a1, a2, a3)


def golden_state(g, stage, net):
    pre = '%s/%s/' % (stage, net)
    return {k[len(pre):]: torch.from_numpy(g[k].copy()) for k in g.files if k.startswith(pre)}


def make_graph(nets, D):
    from b200.graph import Graph
    return Graph(MODEL_2HEAD, nets, {'LSTM_layers': True, 'MLP_layers': False, 'MLP_layers2': False},
                 {'fbank': (0, D)}, {'lab_cd': D, 'lab_mono': D + 1})
