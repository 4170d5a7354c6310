"""Layer spec mirroring the reference FC (mopo/models/fc.py:12-52, :145-154).

Only the description lives here (dims, activation, weight decay, ensemble size); the arithmetic
(fc.py:96-106) is in the CUDA kernels.  Parameters use the reference layout
weights [E,in,out], biases [E,1,out] and init law trunc-normal(std = 1/(2 sqrt(in))), bias 0.
"""
import numpy as np


class FC:
    _activations = (None, 'ReLU', 'tanh', 'sigmoid', 'softmax', 'swish')       # fc.py:15-22

    def __init__(self, output_dim, input_dim=None, activation=None, weight_decay=None, ensemble_size=1):
        if activation not in FC._activations:
            raise ValueError('Invalid activation %r' % (activation,))
        self.input_dim, self.output_dim = input_dim, output_dim
        self.activation = activation
        self.weight_decay = weight_decay
        self.ensemble_size = ensemble_size
        self.weights, self.biases = None, None

    def __repr__(self):                                                         # fc.py:48-52
        return "FC(output_dim={!r}, input_dim={!r}, activation={!r}, weight_decay={!r}, ensemble_size={!r})".format(
            self.output_dim, self.input_dim, self.activation, self.weight_decay, self.ensemble_size)

    def copy(self):
        return FC(self.output_dim, self.input_dim, self.activation, self.weight_decay, self.ensemble_size)

    def get_input_dim(self): return self.input_dim
    def set_input_dim(self, d): self.input_dim = d
    def get_output_dim(self): return self.output_dim
    def set_output_dim(self, d): self.output_dim = d
    def get_activation(self, as_func=True): return self.activation
    def set_activation(self, a): self.activation = a
    def unset_activation(self): self.activation = None
    def get_ensemble_size(self): return self.ensemble_size
    def set_ensemble_size(self, n): self.ensemble_size = n

    def construct_vars(self, rng):
        """fc.py:134-158: trunc-normal(std=1/(2 sqrt(in))) redrawn beyond 2 std; biases zero."""
        if self.input_dim is None or self.output_dim is None:
            raise RuntimeError("Cannot construct variables without fully specifying input and output dimensions.")
        shape = (self.ensemble_size, self.input_dim, self.output_dim)
        w = rng.standard_normal(size=shape)
        bad = np.abs(w) > 2.0
        while bad.any():
            w[bad] = rng.standard_normal(size=int(bad.sum()))
            bad = np.abs(w) > 2.0
        self.weights = (w / (2.0 * np.sqrt(self.input_dim))).astype(np.float32)
        self.biases = np.zeros((self.ensemble_size, 1, self.output_dim), np.float32)
