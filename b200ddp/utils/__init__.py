from .logging import configure_logging, get_logger, RankLogFilter  # noqa: F401
