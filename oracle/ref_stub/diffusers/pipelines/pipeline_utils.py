import contextlib


class _Bar:
    def update(self, n=1):
        pass


class DiffusionPipeline:
    """register_modules / progress_bar / _execution_device / maybe_free_model_hooks as the pipeline's __call__ uses them."""

    def register_modules(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)

    @contextlib.contextmanager
    def progress_bar(self, total=None):
        yield _Bar()

    @property
    def _execution_device(self):
        return self.transformer.device

    def maybe_free_model_hooks(self):
        pass
