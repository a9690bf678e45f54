"""Native library loader (``_native``), GPU solver objects (``gpu``), stand-alone device operators (``operators``)
and the PyTorch extension loader (``torch_ext``)."""
