import logging

from b200ddp.utils.logging import LOG_FORMAT, RankLogFilter


def test_rank_filter_reads_env(monkeypatch):
    rec = logging.LogRecord("x", logging.INFO, __file__, 1, "m", None, None)
    monkeypatch.delenv("RANK", raising=False)
    RankLogFilter().filter(rec)
    assert rec.rank == "?"
    monkeypatch.setenv("RANK", "5")
    RankLogFilter().filter(rec)
    assert rec.rank == "5"


def test_format_is_the_reference_format(reference_module):
    import inspect
    assert LOG_FORMAT in inspect.getsource(reference_module)
