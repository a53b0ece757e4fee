def try_forward(model, x):
    return None


def try_loss(model, x, y, loss_name):
    return None
