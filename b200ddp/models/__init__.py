from .simplenet import SimpleNet  # noqa: F401
