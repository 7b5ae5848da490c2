from logging import *  # noqa: F401,F403
import logging as _l
import datetime as _dt

_cb = []
def set_configure_callback(cb, append=True):
    _cb.append(cb)
def datetime_tag():
    return _dt.datetime.now().strftime("%Y%m%d-%H%M%S")
def configure(*a, **k):
    _l.basicConfig()
def run_main(f, *a, **k):
    return f()
def run_cli(f, *a, **k):
    return f()
def add_file_logger(*a, **k):
    return None
def remove_log_handler(*a, **k):
    return None
def add_memory_logger(*a, **k):
    return None
def get_memory_log():
    return ""
class FileLoggerContext:
    def __init__(self, *a, **k): pass
    def __enter__(self): return self
    def __exit__(self, *a): return False
getLogger = _l.getLogger
Logger = _l.Logger
