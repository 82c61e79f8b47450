"""Optimizers with the exact Keras 2.2.x update rules (epsilon = 1e-7, outside the sqrt).

These objects hold hyper-parameters and, for the CPU oracle backend, the NumPy implementation of the update.
The GPU backend reads the same hyper-parameters into the device ``HyperBlock`` and runs the fused multi-tensor
kernel (csrc/kernels_basic.cuh: optimizer_kernel / csrc/allreduce.cuh: allreduce_opt_kernel).

Reference: optimizer names at mnist.py:45,57, DistHPO_mnist.ipynb:145; ``getattr(optimizers, name)(lr=lr)`` at
rpv.py:62; ``Adadelta(1.0 * hvd.size())`` at DistTrain_mnist.ipynb:307.
"""
import numpy as np

EPSILON = 1e-7


class Optimizer(object):
    kind = -1            # id understood by the CUDA kernels: 0 SGD, 1 Adam, 2 Nadam, 3 Adadelta
    n_slots = 0

    def __init__(self, lr, decay=0.0):
        self.lr = float(lr)
        self.decay = float(decay)
        self.iterations = 0

    # ---- oracle (NumPy) implementation -------------------------------------------------
    def init_slots(self, params):
        return [[np.zeros_like(p) for p in params] for _ in range(self.n_slots)]

    def current_lr(self):
        lr = self.lr
        if self.decay > 0:
            lr = lr * (1.0 / (1.0 + self.decay * self.iterations))
        return lr

    def apply(self, params, grads, slots):
        """In-place update of ``params`` (list of float32 arrays); returns nothing."""
        raise NotImplementedError

    # ---- (de)serialisation, Keras training_config style ----------------------------------
    def get_config(self):
        return {"lr": float(self.lr), "decay": float(self.decay)}

    @property
    def class_name(self):
        return type(self).__name__

    def hyper(self):
        """Fields of the device HyperBlock."""
        return dict(lr=self.lr, beta1=0.9, beta2=0.999, eps=EPSILON, rho=0.95, momentum=0.0,
                    schedule_decay=0.004, decay=self.decay, m_schedule=1.0, iterations=self.iterations)

    # Keras `optimizer.weights` order, for /optimizer_weights in checkpoints
    def state_arrays(self, slots):
        out = [np.asarray(self.iterations, dtype=np.int64)]
        for s in slots:
            out.extend(s)
        return out

    def load_state_arrays(self, arrays, slots):
        arrays = list(arrays)
        self.iterations = int(np.asarray(arrays.pop(0)).reshape(-1)[0])
        for s in slots:
            for i in range(len(s)):
                s[i][...] = arrays.pop(0)


class SGD(Optimizer):
    kind = 0
    n_slots = 1

    def __init__(self, lr=0.01, momentum=0.0, decay=0.0, nesterov=False):
        super(SGD, self).__init__(lr, decay)
        self.momentum = float(momentum)
        if nesterov:
            raise NotImplementedError("nesterov momentum is not used by the reference")

    def apply(self, params, grads, slots):
        lr = self.current_lr()
        for p, g, m in zip(params, grads, slots[0]):
            m[...] = self.momentum * m - lr * g
            p += m
        self.iterations += 1

    def get_config(self):
        return {"lr": self.lr, "momentum": self.momentum, "decay": self.decay, "nesterov": False}

    def hyper(self):
        h = super(SGD, self).hyper()
        h["momentum"] = self.momentum
        return h


class Adam(Optimizer):
    kind = 1
    n_slots = 2

    def __init__(self, lr=0.001, beta_1=0.9, beta_2=0.999, epsilon=None, decay=0.0, amsgrad=False):
        super(Adam, self).__init__(lr, decay)
        self.beta_1, self.beta_2 = float(beta_1), float(beta_2)
        self.epsilon = EPSILON if epsilon is None else float(epsilon)
        if amsgrad:
            raise NotImplementedError("amsgrad is not used by the reference")

    def apply(self, params, grads, slots):
        lr = self.current_lr()
        t = self.iterations + 1
        lr_t = np.float32(lr * np.sqrt(1.0 - self.beta_2 ** t) / (1.0 - self.beta_1 ** t))
        b1, b2, eps = np.float32(self.beta_1), np.float32(self.beta_2), np.float32(self.epsilon)
        for p, g, m, v in zip(params, grads, slots[0], slots[1]):
            m[...] = b1 * m + (1 - b1) * g
            v[...] = b2 * v + (1 - b2) * g * g
            p -= lr_t * m / (np.sqrt(v) + eps)
        self.iterations += 1

    def get_config(self):
        return {"lr": self.lr, "beta_1": self.beta_1, "beta_2": self.beta_2, "decay": self.decay,
                "epsilon": self.epsilon, "amsgrad": False}

    def hyper(self):
        h = super(Adam, self).hyper()
        h.update(beta1=self.beta_1, beta2=self.beta_2, eps=self.epsilon)
        return h


class Nadam(Optimizer):
    kind = 2
    n_slots = 2

    def __init__(self, lr=0.002, beta_1=0.9, beta_2=0.999, epsilon=None, schedule_decay=0.004):
        super(Nadam, self).__init__(lr, 0.0)
        self.beta_1, self.beta_2 = float(beta_1), float(beta_2)
        self.epsilon = EPSILON if epsilon is None else float(epsilon)
        self.schedule_decay = float(schedule_decay)
        self.m_schedule = 1.0

    def apply(self, params, grads, slots):
        t = self.iterations + 1
        mu_t = self.beta_1 * (1.0 - 0.5 * 0.96 ** (t * self.schedule_decay))
        mu_t1 = self.beta_1 * (1.0 - 0.5 * 0.96 ** ((t + 1) * self.schedule_decay))
        sched_new = self.m_schedule * mu_t
        sched_next = sched_new * mu_t1
        self.m_schedule = sched_new
        b1, b2, eps = np.float32(self.beta_1), np.float32(self.beta_2), np.float32(self.epsilon)
        for p, g, m, v in zip(params, grads, slots[0], slots[1]):
            g_prime = g / np.float32(1.0 - sched_new)
            m[...] = b1 * m + (1 - b1) * g
            m_prime = m / np.float32(1.0 - sched_next)
            v[...] = b2 * v + (1 - b2) * g * g
            v_prime = v / np.float32(1.0 - self.beta_2 ** t)
            m_bar = np.float32(1.0 - mu_t) * g_prime + np.float32(mu_t1) * m_prime
            p -= np.float32(self.lr) * m_bar / (np.sqrt(v_prime) + eps)
        self.iterations += 1

    def get_config(self):
        return {"lr": self.lr, "beta_1": self.beta_1, "beta_2": self.beta_2, "epsilon": self.epsilon,
                "schedule_decay": self.schedule_decay}

    def hyper(self):
        h = super(Nadam, self).hyper()
        h.update(beta1=self.beta_1, beta2=self.beta_2, eps=self.epsilon, schedule_decay=self.schedule_decay,
                 m_schedule=self.m_schedule)
        return h

    def state_arrays(self, slots):
        out = [np.asarray(self.iterations, dtype=np.int64), np.asarray(self.m_schedule, dtype=np.float32)]
        for s in slots:
            out.extend(s)
        return out

    def schedule_after(self, iterations):
        """The running product m_schedule after `iterations` steps (Keras keeps it as a variable but does NOT save it)."""
        prod = 1.0
        for t in range(1, int(iterations) + 1):
            prod *= self.beta_1 * (1.0 - 0.5 * 0.96 ** (t * self.schedule_decay))
        return prod

    def load_state_arrays(self, arrays, slots):
        """Two layouts: ours [iterations, m_schedule] + ms + vs, and Keras 2.2's [iterations] + ms + vs (its Nadam.weights has no
        m_schedule); told apart by the array count.  For the Keras layout the schedule is rebuilt from the step count.  Nothing is
        modified unless the whole list fits."""
        arrays = list(arrays)
        n = sum(len(s) for s in slots)
        if len(arrays) not in (1 + n, 2 + n):
            raise ValueError("Nadam state: %d arrays, expected %d or %d" % (len(arrays), 1 + n, 2 + n))
        tensors = arrays[len(arrays) - n:]
        flat = [t for s in slots for t in s]
        for dst, src in zip(flat, tensors):
            if np.shape(dst) != np.shape(src):
                raise ValueError("Nadam state: shape mismatch %r vs %r" % (np.shape(src), np.shape(dst)))
        iterations = int(np.asarray(arrays[0]).reshape(-1)[0])
        self.m_schedule = float(np.asarray(arrays[1]).reshape(-1)[0]) if len(arrays) == 2 + n else self.schedule_after(iterations)
        self.iterations = iterations
        for dst, src in zip(flat, tensors):
            dst[...] = src


class Adadelta(Optimizer):
    kind = 3
    n_slots = 2

    def __init__(self, lr=1.0, rho=0.95, epsilon=None, decay=0.0):
        super(Adadelta, self).__init__(lr, decay)
        self.rho = float(rho)
        self.epsilon = EPSILON if epsilon is None else float(epsilon)

    def apply(self, params, grads, slots):
        lr = np.float32(self.current_lr())
        rho, eps = np.float32(self.rho), np.float32(self.epsilon)
        for p, g, a, d in zip(params, grads, slots[0], slots[1]):
            a[...] = rho * a + (1 - rho) * g * g
            u = g * np.sqrt(d + eps) / np.sqrt(a + eps)
            p -= lr * u
            d[...] = rho * d + (1 - rho) * u * u
        self.iterations += 1

    def get_config(self):
        return {"lr": self.lr, "rho": self.rho, "decay": self.decay, "epsilon": self.epsilon}

    def hyper(self):
        h = super(Adadelta, self).hyper()
        h.update(rho=self.rho, eps=self.epsilon)
        return h


_BY_NAME = {"sgd": SGD, "adam": Adam, "nadam": Nadam, "adadelta": Adadelta}


def get(identifier, **kwargs):
    """Keras-style lookup: a name ('Adam'), a class, an instance, or a {'class_name','config'} dict."""
    if isinstance(identifier, Optimizer):
        return identifier
    if isinstance(identifier, dict):
        return get(identifier["class_name"], **identifier.get("config", {}))
    if isinstance(identifier, type) and issubclass(identifier, Optimizer):
        return identifier(**kwargs)
    key = str(identifier).lower()
    if key not in _BY_NAME:
        raise ValueError("unknown optimizer %r (have: SGD, Adam, Nadam, Adadelta)" % (identifier,))
    kwargs.pop("amsgrad", None) if key != "adam" else None
    return _BY_NAME[key](**kwargs)
