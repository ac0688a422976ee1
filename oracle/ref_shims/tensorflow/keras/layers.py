from ._core import _InputLayer as InputLayer, _Dense as Dense
Input = InputLayer
