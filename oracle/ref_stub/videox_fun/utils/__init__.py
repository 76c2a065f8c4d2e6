def cfg_skip():
    def deco(fn):
        return fn
    return deco
