class AutoencoderKLWan:      # type annotations of the pipeline's ctor only; oracle/ref_pipeline.py provides the fakes
    pass


class AutoTokenizer:
    pass


class WanT5EncoderModel:
    pass
