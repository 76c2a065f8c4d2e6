class FlowUniPCMultistepScheduler:
    """Name placeholder for the isinstance() branch at pipeline_wan_versecrafter.py:750; VideoX-Fun is un-vendored.  The
    in-repo implementation of this sampler is versecrafter_b200.schedulers.FlowUniPCMultistepScheduler (GPU kernels)."""
