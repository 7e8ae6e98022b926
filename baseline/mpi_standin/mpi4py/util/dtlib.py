"""mpi4py.util.dtlib.to_numpy_dtype for the stand-in (reference ddl/connection.py:105-106)."""


def to_numpy_dtype(datatype):
    return datatype.np_dtype
