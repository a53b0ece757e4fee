def native_available(plan, x):
    return False


def trunk(plan, x):
    return None
