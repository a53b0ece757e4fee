def native_available(x, units):
    return False
