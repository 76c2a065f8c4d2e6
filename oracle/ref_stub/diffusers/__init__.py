"""Import stub (test infrastructure): just enough of `diffusers` for the reference's model files and its sampling pipeline
(versecrafter/pipeline/pipeline_wan_versecrafter.py) to import and run in a container that does not have diffusers
installed.  The only arithmetic here is the input pre-processing of the fake image processor and the Euler scheduler of
oracle/ref_pipeline.py; the transformer forward and the loop body are the reference's own code."""
from .schedulers import FlowMatchEulerDiscreteScheduler  # noqa: F401


class AutoencoderKL:  # imported by name only
    pass
