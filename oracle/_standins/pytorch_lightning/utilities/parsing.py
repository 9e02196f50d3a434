class AttributeDict(dict):
    """dict with attribute access; the only pytorch_lightning symbol the reference path uses."""

    def __getattr__(self, key):
        try:
            return self[key]
        except KeyError as exc:
            raise AttributeError(key) from exc

    def __setattr__(self, key, value):
        self[key] = value
