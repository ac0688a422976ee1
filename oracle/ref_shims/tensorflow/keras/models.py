from ._core import Sequential


def load_model(path, compile=False):
    raise NotImplementedError("tf-shim: SavedModel loading is not restated")
