"""Context: device / dtype / backend policy.  Mirrors `lettuce.Context`
(reference/lettuce/_context.py:9-107): same constructor defaults and tensor factories."""
from typing import List, Optional, Union

import numpy as np
import torch

__all__ = ['Context']


class Context:
    def __init__(self, device: Optional[Union[torch.device, str]] = None,
                 dtype: Optional[torch.dtype] = None, use_native: Optional[bool] = None):
        cuda = torch.cuda.is_available()
        if device is None and use_native is None:          # _context.py:17-20
            device = 'cuda:0' if cuda else 'cpu'
            use_native = cuda
        elif device is None:                                # :22-26
            assert not use_native or cuda, \
                'cuda_native extension explicitly requested but cuda is not available!'
            device = 'cuda:0'
        else:
            is_cuda = 'cuda' in str(device)
            if is_cuda:
                assert cuda, 'cuda device explicitly requested but cuda is not available!'
            else:
                assert 'cpu' in str(device), \
                    (f"lettuce is designed to work on cpu or cuda devices. "
                     f"{device} is not supported!")
            if use_native is None:                          # :28-38
                use_native = is_cuda
            else:                                           # :40-51
                assert is_cuda or not use_native, \
                    ('can not use explicitly requested cuda_native extension on '
                     'explicitly requested cpu device!')
        dtype = dtype or torch.float32
        assert dtype in [torch.float16, torch.float32, torch.float64], \
            f"lettuce is designed to work with common float types (16, 32 and 64 bit). {dtype} is not supported!"
        self.device = torch.device(device)
        self.dtype = dtype
        self.use_native = use_native

    def empty_tensor(self, size, *args, dtype=None, **kwargs) -> torch.Tensor:
        return torch.empty(size, *args, **kwargs, device=self.device, dtype=(dtype or self.dtype))

    def zero_tensor(self, size, *args, dtype=None, **kwargs) -> torch.Tensor:
        return torch.zeros(size, *args, **kwargs, device=self.device, dtype=(dtype or self.dtype))

    def one_tensor(self, size, *args, dtype=None, **kwargs) -> torch.Tensor:
        return torch.ones(size, *args, **kwargs, device=self.device, dtype=(dtype or self.dtype))

    def convert_to_tensor(self, array, *args, dtype: Optional[torch.dtype] = None, **kwargs) -> torch.Tensor:
        """bool stays bool, uint8 stays uint8, everything else becomes the context dtype
        (_context.py:79-100)."""
        new_dtype = dtype
        if new_dtype is None:
            adt = getattr(array, 'dtype', None)
            if adt in (bool, torch.bool, np.bool_):
                new_dtype = torch.bool
            elif adt in (torch.uint8, np.uint8):
                new_dtype = torch.uint8
            else:
                new_dtype = self.dtype
        if isinstance(array, torch.Tensor):
            return array.to(*args, **kwargs, device=self.device, dtype=new_dtype)
        return torch.tensor(array, *args, **kwargs, device=self.device, dtype=new_dtype)

    @staticmethod
    def convert_to_ndarray(tensor: Union[torch.Tensor, List]) -> np.ndarray:
        if isinstance(tensor, torch.Tensor):
            return tensor.detach().cpu().numpy()
        return np.array(tensor)
