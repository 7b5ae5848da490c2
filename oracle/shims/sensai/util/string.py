class ToStringMixin:
    def _tostring_excludes(self): return []
    def _tostring_includes(self): return []
    def _tostring_additional_entries(self): return {}
    def _tostring_exclude_private(self): return False
    def _tostring_exclude_exceptions(self): return []
    def _tostring_object_info(self): return ""
    def __repr__(self):
        return f"{self.__class__.__name__}[...]"
    __str__ = __repr__
    def pprint(self, *a, **k): print(repr(self))
    def pprints(self, *a, **k): return repr(self)
def pretty_string_repr(o, *a, **k): return repr(o)
def list_string(l, *a, **k): return str(list(l))
def dict_string(d, *a, **k): return str(dict(d))
