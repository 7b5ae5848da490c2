import pickle
def setstate(cls, obj, state, renamed_properties=None, new_optional_properties=None,
             new_default_properties=None, removed_properties=None):
    if renamed_properties:
        for old, new in renamed_properties.items():
            if old in state:
                state[new] = state.pop(old)
    if new_optional_properties:
        for p in new_optional_properties:
            state.setdefault(p, None)
    if new_default_properties:
        for p, v in new_default_properties.items():
            state.setdefault(p, v)
    if removed_properties:
        for p in removed_properties:
            state.pop(p, None)
    parent = super(cls, obj)
    if hasattr(parent, "__setstate__") and parent.__setstate__ is not None:
        try:
            parent.__setstate__(state)
            return
        except Exception:
            pass
    obj.__dict__.update(state)
def dump_pickle(obj, path, **_k):
    with open(path, "wb") as f:
        pickle.dump(obj, f)
def load_pickle(path, **_k):
    with open(path, "rb") as f:
        return pickle.load(f)
