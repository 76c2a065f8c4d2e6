class SimpleAdapter:
    def __init__(self, *a, **k):
        raise RuntimeError("oracle stub: control adapter unused by VerseCrafter")
