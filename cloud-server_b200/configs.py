"""Canonical net_config values (the JSON `cnn()` interprets, construct_distribute.py:208-265)."""

# MNIST LeNet in the reference DSL (the commented net in construct_distribute_url.py:307-334):
# conv5x5x32+b, relu, pool, conv5x5x64+b, relu, pool, connect 1024, relu, (implicit output 10)
LENET = {"middle_layer": [
    {"layer": "conv", "filter": [5, 5, 32], "isBias": "True"}, {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
    {"layer": "conv", "filter": [5, 5, 64], "isBias": "True"}, {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
    {"layer": "connect", "hidden": 1024}, {"layer": "active", "active_func": "relu"}]}

# BASELINE config 5: 4 conv + 3 dense on 32x32x3
DEEPNET = {"middle_layer": [
    {"layer": "conv", "filter": [3, 3, 32], "isBias": "True"}, {"layer": "active", "active_func": "relu"},
    {"layer": "conv", "filter": [3, 3, 32], "isBias": "True"}, {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
    {"layer": "conv", "filter": [3, 3, 64], "isBias": "True"}, {"layer": "active", "active_func": "relu"},
    {"layer": "conv", "filter": [3, 3, 64], "isBias": "True"}, {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
    {"layer": "connect", "hidden": 512}, {"layer": "active", "active_func": "relu"},
    {"layer": "connect", "hidden": 256}, {"layer": "active", "active_func": "relu"}]}

# the example of API.md:306-332 (defaults: no conv bias, sigmoid, hidden 512)
API_EXAMPLE = {"middle_layer": [
    {"layer": "conv", "filter": [2, 2, 10]}, {"layer": "conv", "filter": [2, 2, 20]}, {"layer": "pool"},
    {"layer": "norm"}, {"layer": "active"}, {"layer": "connect"}, {"layer": "connect"}]}
