from typing import Optional, Tuple, Union
from torch import Tensor


class SparseTensor:  # never instantiated in the shimmed path
    pass


Adj = Union[Tensor, SparseTensor]
OptTensor = Optional[Tensor]
PairTensor = Tuple[Tensor, Tensor]
