"""Logger shim: loguru when installed (the reference's choice, dist.py:6), stdlib otherwise."""
try:  # pragma: no cover - trivial
    from loguru import logger  # type: ignore
except Exception:  # pragma: no cover
    import logging as _logging

    logger = _logging.getLogger("llmss_b200")
    if not logger.handlers:
        _h = _logging.StreamHandler()
        _h.setFormatter(_logging.Formatter("%(asctime)s %(levelname)s %(message)s"))
        logger.addHandler(_h)
        logger.setLevel(_logging.INFO)
