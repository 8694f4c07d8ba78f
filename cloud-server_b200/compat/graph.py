"""The reference's graph-building vocabulary (construct_distribute.py:57-311) over the CUDA library.

In the reference these functions append TensorFlow ops; here they append layer records to a
symbolic tensor, and ``compile_net`` hands the resulting ``middle_layer`` list to nc_net_create.
Names, argument order and defaults are the reference's, so code written against it keeps working:

    x = placeholder(28, 28, 1)
    h = conv2d(x, [5, 5, 32], isBias=True); h = active(h, "relu"); h = max_pool(h)
    y = cnn({"middle_layer": [...]}, x)               # or the JSON interpreter
    net = compile_net(y)
"""
import json


class Sym:
    """A symbolic activation: the layer list that produces it + its static shape."""

    def __init__(self, layers, shape, input_shape):
        self.layers, self.shape, self.input_shape = layers, tuple(shape), tuple(input_shape)

    def get_shape(self):
        return (None,) + self.shape

    def _then(self, layer, shape):
        return Sym(self.layers + [layer], shape, self.input_shape)


def placeholder(h=28, w=28, c=1):
    """tf.placeholder(tf.float32, [None, 784]) + tf.reshape(x, [-1, 28, 28, 1]) (:216, :362)."""
    return Sym([], (h, w, c), (h, w, c))


def _out(n, k, s, padding):
    return -(-n // s) if padding == "SAME" else -(-(n - k + 1) // s)


def weight_variable(shape, init="norm", stddev_norm=0.1):                       # :57-76
    if init not in ("zero", "norm", "xavier"):
        assert False, init + " is not a legal init"
    return {"shape": list(shape), "init": init, "stddev_norm": stddev_norm}


def bias_variable(shape, constant=0.1):                                         # :79-87
    return {"shape": list(shape), "constant": constant}


def conv2d(inputs, filters, stride=(1, 1), padding='SAME', init="norm", isBias=False, bias_constant=0.1,
           stddev_norm=0.1):                                                    # :91-118
    weight_variable([filters[0], filters[1], inputs.shape[-1], filters[2]], init=init, stddev_norm=stddev_norm)
    h, w, _ = inputs.shape
    layer = {"layer": "conv", "filter": list(filters), "stride": list(stride), "padding": padding, "init": init,
             "isBias": "True" if isBias else "False", "bias_constant": bias_constant, "stddev_norm": stddev_norm}
    return inputs._then(layer, (_out(h, filters[0], stride[0], padding), _out(w, filters[1], stride[1], padding), filters[2]))


def max_pool(x, kernel=(2, 2), stride=(2, 2), padding="SAME"):                  # :121-130
    h, w, c = x.shape
    layer = {"layer": "pool", "kernel": list(kernel), "stride": list(stride), "padding": padding}
    return x._then(layer, (_out(h, kernel[0], stride[0], padding), _out(w, kernel[1], stride[1], padding), c))


def active(x, active_func, param=None):                                         # :133-152
    if param is None:
        param = [0.2]
    if active_func not in ("relu", "sigmoid", "leaky_relu"):
        assert False, active_func + " is not a legal activation"
    return x._then({"layer": "active", "active_func": active_func, "param": list(param)}, x.shape)


def normalize(inputs, var_epsilon=1e-3):                                        # :155-165
    return inputs._then({"layer": "norm", "epsilon": var_epsilon}, inputs.shape)


def connect_layer(x, hidden):                                                   # :168-182
    return x._then({"layer": "connect", "hidden": hidden}, (hidden,))


def cnn(net_config, x):                                                         # :208-265
    """The JSON interpreter; the implicit output layer (:253-264) is added by the library."""
    if isinstance(net_config, str):
        net_config = json.loads(net_config)
    cur = x
    for L in net_config["middle_layer"]:
        name = L["layer"]
        if name == "conv":
            cur = conv2d(cur, L["filter"], L.get("stride", [1, 1]), L.get("padding", "SAME"), init=L.get("init", "norm"),
                         isBias=False if "isBias" not in L else L["isBias"] != "False",
                         bias_constant=L.get("bias_constant", 0.1), stddev_norm=L.get("stddev_norm", 0.1))
        elif name == "pool":
            cur = max_pool(cur, L.get("kernel", [2, 2]), L.get("stride", [2, 2]), L.get("padding", "SAME"))
        elif name == "active":
            cur = active(cur, L.get("active_func", "sigmoid"), L.get("param", [0.2]))
        elif name == "connect":
            cur = connect_layer(cur, L.get("hidden", 512))
        elif name == "norm":
            cur = normalize(cur, L.get("epsilon", 1e-3))
    return Sym(cur.layers, (10,), cur.input_shape)


def get_net(net_type, net_config, x, keep_prob=None):                           # :269-281 (net_type ignored there too)
    return cnn(net_config, x)


def loss_function(name, logits, labels=None):                                   # :285-298
    if name == "mse" or name == "entropy":
        return name
    return None          # the reference returns None here and then crashes in minimize()


def optimizer_function(optimizer_name, learning_rate):                          # :301-311
    if optimizer_name == "GradientDescentOptimizer":
        return ("sgd", float(learning_rate))
    return ("adagrad", float(learning_rate))


def net_config_of(sym):
    return {"middle_layer": list(sym.layers)}


def compile_net(sym, ctx=None, classes=10, seed=1):
    from cloud_server_b200.net import Net
    h, w, c = sym.input_shape
    return Net(net_config_of(sym), ctx, h, w, c, classes, seed)
