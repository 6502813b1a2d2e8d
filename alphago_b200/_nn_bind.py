"""ctypes signatures of the network entry points (stage 2); filled in as they are built."""


def bind(L):
    return L
