def get_1d_rotary_pos_embed(*a, **k):
    raise RuntimeError("oracle stub: unused by the VerseCrafter pipeline")
