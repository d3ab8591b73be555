"""Host-layer logic that needs no GPU: datum -> sfv conversion and config validation."""
import json

import pytest

from jubatus_b200 import nn

CONVERTER = {
    "string_filter_types": {}, "string_filter_rules": [], "num_filter_types": {}, "num_filter_rules": [],
    "string_types": {}, "string_rules": [{"key": "*", "type": "str", "sample_weight": "bin", "global_weight": "bin"}],
    "num_types": {}, "num_rules": [{"key": "*", "type": "num"}],
}


def test_num_and_str_rules():
    d = nn.Datum().add_number("age", 31).add_string("name", "juba").add_number("w", -0.5)
    sfv = nn.convert(json.dumps(CONVERTER), d)
    # strings first, then numbers, in datum order (datum_to_fv_converter::convert)
    assert sfv == [("name$juba@str#bin/bin", 1.0), ("age@num", 31.0), ("w@num", -0.5)]


def test_key_matchers_and_log():
    conv = dict(CONVERTER, num_rules=[{"key": "a*", "type": "num"}, {"key": "*z", "type": "log"}], string_rules=[])
    d = nn.Datum().add_number("ab", 2).add_number("xz", 100).add_number("q", 1)
    sfv = nn.convert(json.dumps(conv), d)
    assert sfv[0] == ("ab@num", 2.0)
    assert sfv[1][0] == "xz@log" and abs(sfv[1][1] - 4.60517) < 1e-5
    assert len(sfv) == 2


def test_converter_errors():
    with pytest.raises(nn.JubatusError):
        nn.convert("{not json", nn.Datum())
    with pytest.raises(nn.JubatusError):
        nn.convert(json.dumps(dict(CONVERTER, bogus=1)), nn.Datum())
    with pytest.raises(nn.JubatusError):
        nn.convert(json.dumps(dict(CONVERTER, num_rules=[{"key": "*", "type": "nope"}])), nn.Datum())


def test_shipped_config_schema_is_accepted_by_the_parser():
    # the reference's example configs (config/nearest_neighbor/*.json) share this shape
    cfg = {"method": "lsh", "parameter": {"hash_num": 64}, "converter": CONVERTER}
    sfv = nn.convert(json.dumps(cfg["converter"]), nn.Datum().add_number("key", 1.0))
    assert sfv == [("key@num", 1.0)]


def test_unlearner_policies():
    """core::unlearner lru / random (config/*_unlearn_lru.json): the policy alone."""
    ev, size = nn.unlearner_trace("lru", {"max_size": 3}, ["a", "b", "c", "a", "d", "e", "-a", "f", "g"])
    # a b c | touch a | d evicts b (least recently used) | e evicts c | remove a | f fits | g evicts d
    assert ev == [(4, "b"), (5, "c"), (8, "d")] and size == 3
    ev, size = nn.unlearner_trace("lru", {"max_size": 2}, ["x", "x", "x", "y", "x", "z"])
    assert ev == [(5, "y")] and size == 2
    ids = [str(i) for i in range(50)]
    ev1, size1 = nn.unlearner_trace("random", {"max_size": 5, "seed": 7}, ids)
    ev2, _ = nn.unlearner_trace("random", {"max_size": 5, "seed": 7}, ids)
    ev3, _ = nn.unlearner_trace("random", {"max_size": 5, "seed": 8}, ids)
    assert size1 == 5 and len(ev1) == 45 and ev1 == ev2 and ev1 != ev3
    alive = set()
    for pos, i in enumerate(ids):            # a victim is always a row that is held at that moment
        for p, v in ev1:
            if p == pos:
                assert v in alive
                alive.discard(v)
        alive.add(i)
    assert len(alive) == 5
    for name, param in (("lru", {"max_size": 0}), ("lru", {}), ("fifo", {"max_size": 3}), ("lru", {"max_size": 3, "bogus": 1}),
                        ("lru", {"max_size": "3"})):
        with pytest.raises(nn.JubatusError):
            nn.unlearner_trace(name, param, ["a"])


def test_combination_rules_and_unsupported_sections():
    """config/anomaly/light_lof_combinational_feature.json:15-18: pairwise "mul" / "add" features over
    the converted features; sections without an implementation must be empty (a silently ignored
    rule would change every signature)."""
    conv = dict(CONVERTER, combination_types={}, combination_rules=[{"key_left": "*", "key_right": "*", "type": "mul"}])
    d = nn.Datum().add_string("s", "x").add_number("a", 2).add_number("b", 3)
    sfv = nn.convert(json.dumps(conv), d)
    assert sfv[:3] == [("s$x@str#bin/bin", 1.0), ("a@num", 2.0), ("b@num", 3.0)]
    assert sfv[3:] == [("s$x@str#bin/bin&a@num/mul", 2.0), ("s$x@str#bin/bin&b@num/mul", 3.0), ("a@num&b@num/mul", 6.0)]
    conv2 = dict(conv, combination_rules=[{"key_left": "a*", "key_right": "b*", "type": "add"}])
    assert nn.convert(json.dumps(conv2), d)[3:] == [("a@num&b@num/add", 5.0)]
    conv3 = dict(conv, combination_rules=[{"key_left": "b*", "key_right": "a*", "type": "add"}])   # either order matches
    assert nn.convert(json.dumps(conv3), d)[3:] == [("a@num&b@num/add", 5.0)]
    with pytest.raises(nn.JubatusError):
        nn.convert(json.dumps(dict(conv, combination_rules=[{"key_left": "*", "key_right": "*", "type": "div"}])), d)
    for section, value in (("num_filter_rules", [{"key": "*", "type": "f", "suffix": "_"}]), ("string_types", {"bigram": {"method": "ngram"}}),
                           ("binary_rules", [{"key": "*", "type": "b"}]), ("combination_types", {"c": {"method": "dynamic"}})):
        with pytest.raises(nn.JubatusError, match="not supported"):
            nn.convert(json.dumps(dict(conv, **{section: value})), d)
