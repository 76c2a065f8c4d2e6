def get_sequence_parallel_rank():
    return 0


def get_sequence_parallel_world_size():
    return 1


def get_sp_group():
    raise RuntimeError("oracle stub: no sequence-parallel group (world size 1)")


def usp_attn_forward(*a, **k):
    raise RuntimeError("oracle stub: usp_attn_forward is not available")


class xFuserLongContextAttention:
    pass
