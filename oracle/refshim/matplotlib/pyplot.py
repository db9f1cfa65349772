"""stub: the reference only plots after the run"""


def __getattr__(name):
    def _noop(*a, **kw):
        return None
    return _noop
