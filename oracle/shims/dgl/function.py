import torch


class BinaryMessage:
    def __init__(self, kind, op, lhs, rhs, out):
        self.kind, self.op, self.lhs, self.rhs, self.out = kind, op, lhs, rhs, out


class Reduce:
    def __init__(self, name, msg, out):
        self.name, self.msg, self.out = name, msg, out


def u_add_v(lhs, rhs, out):
    return BinaryMessage("u_op_v", torch.add, lhs, rhs, out)


def u_mul_e(lhs, rhs, out):
    return BinaryMessage("u_op_e", torch.mul, lhs, rhs, out)


def sum(msg, out):  # noqa: A001
    return Reduce("sum", msg, out)
