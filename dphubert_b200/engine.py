"""placeholder — replaced below"""


def fused_num_params(model):
    return None


def invalidate(model):
    pass
