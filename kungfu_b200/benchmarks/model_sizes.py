"""Gradient-size tables of the benchmark models, derived from the model
definitions of this repo (the reference ships literal tables,
/root/reference/srcs/python/kungfu/tensorflow/v1/benchmarks/model_sizes.py:1-81)."""


def _sizes(model):
    return [p.numel() for p in model.parameters() if p.requires_grad]


def model_sizes(name):
    from .. import models
    name = name.lower()
    if name == 'resnet50':
        return _sizes(models.resnet50())
    if name == 'vgg16':
        return _sizes(models.vgg16())
    if name == 'densenet121':
        return _sizes(models.densenet121())
    if name == 'bert':
        # parameters + two Adam slots per parameter, like the reference's list
        s = _sizes(models.bert_base())
        return s + s + s
    raise ValueError('unknown model %s' % name)
