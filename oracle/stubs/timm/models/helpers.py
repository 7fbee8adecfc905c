def build_model_with_cfg(*a, **k):
    raise NotImplementedError

def named_apply(fn, module, name='', depth_first=True, include_root=False):
    raise NotImplementedError

def adapt_input_conv(*a, **k):
    raise NotImplementedError
