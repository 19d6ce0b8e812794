"""tcgen05 GEMM bindings (filled in with csrc/gemm_tc.cu)."""


def conv_forward(x, w, bias, y, shp) -> bool:
    return False


def conv_dgrad(dy, w, dx, shp) -> bool:
    return False


def conv_wgrad(x, dy, dw, db, shp) -> bool:
    return False
