"""Weight / mask dumps written on the first valid/forward call (reference: save_cgs_mat.py:7-20)."""
import os

import scipy.io as sio


def _dump(t, path, key):
    sio.savemat(path, {key: t.detach().cpu().numpy()})
    return 1


def save_mat(in_mat, mat_num='1', dir='/home/dkadetot/pytorch-kaldi/saved_mat'):
    if not os.path.isdir(dir):
        os.mkdir(dir)
    return _dump(in_mat, dir + '/mat%s.mat' % mat_num, 'M%s' % mat_num)


def save_hcgs_mat(in_mat, mat_num='1', dir='/home/dkadetot/pytorch-kaldi/saved_mat'):
    if not os.path.isdir(dir):
        os.mkdir(dir)
    return _dump(in_mat, dir + '/conn_mat%s.mat' % mat_num, 'CM%s' % mat_num)
