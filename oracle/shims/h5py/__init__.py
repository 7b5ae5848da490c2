class Group: pass
class File(Group):
    def __init__(self, *a, **k): raise RuntimeError("h5py shim: HDF5 not available")
class Dataset: pass
