def resize_pos_embed(*a, **k):
    raise NotImplementedError
