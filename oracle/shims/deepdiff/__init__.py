class DeepDiff(dict):
    """Shim: empty dict == 'no difference'. Only used by Batch.__eq__ in the reference."""
    def __init__(self, a, b, **_k):
        super().__init__()
        import pickle
        try:
            if pickle.dumps(a) != pickle.dumps(b):
                self["values_changed"] = True
        except Exception:
            if a != b:
                self["values_changed"] = True
