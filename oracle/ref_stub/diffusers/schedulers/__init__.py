class FlowMatchEulerDiscreteScheduler:
    """Name placeholder: the pipeline only does isinstance() checks against it (pipeline_wan_versecrafter.py:746).  The
    scheduler the boundary test passes (oracle/ref_pipeline.EulerFlowScheduler) subclasses it."""
    order = 1
