"""``pytorch_lightning.loggers.logger.Logger`` stand-in (see package docstring)."""


class Logger:
    def __init__(self, *a, **kw) -> None:
        pass
