"""AttributeDict: the one pytorch_lightning symbol the reference's configs use
(planner_factory.py:2, utils/universal_factory.py:3).  Lightning's own class is used when importable."""
try:  # pragma: no cover - lightning is not in this image
    from pytorch_lightning.utilities.parsing import AttributeDict  # type: ignore
except Exception:
    class AttributeDict(dict):
        def __getattr__(self, key):
            try:
                return self[key]
            except KeyError as exc:
                raise AttributeError(f"Missing attribute '{key}'") from exc

        def __setattr__(self, key, value):
            self[key] = value
