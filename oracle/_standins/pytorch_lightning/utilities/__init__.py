from .parsing import AttributeDict  # noqa: F401
