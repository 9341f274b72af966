"""Stand-in for `h5py` (absent here), used ONLY by tests/golden/make_golden_io.py so that the unmodified
reference `export_predictions` runs: groups and datasets are kept in memory (`h5py.LAST_FILE`)."""
import numpy as np

LAST_FILE = None


class _Group(dict):
    def create_dataset(self, name, data=None):
        self[name] = np.array(data)
        return self[name]


class File(dict):
    def __init__(self, path, mode="r"):
        super().__init__()
        global LAST_FILE
        self.path, self.mode = path, mode
        LAST_FILE = self

    def create_group(self, name):
        if name in self:
            raise RuntimeError(f"group {name} exists")
        self[name] = _Group()
        return self[name]

    def close(self):
        pass
