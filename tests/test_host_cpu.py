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
    for section, value in (("num_filter_rules", [{"key": "*", "type": "f", "suffix": "_"}]), ("string_types", {"morph": {"method": "dynamic", "path": "libmecab_splitter.so"}}),
                           ("binary_rules", [{"key": "*", "type": "b"}]), ("combination_types", {"c": {"method": "dynamic"}})):
        with pytest.raises(nn.JubatusError, match="not supported"):
            nn.convert(json.dumps(dict(conv, **{section: value})), d)


def test_converter_sections_of_the_shipped_configs_parse():
    """Every config the reference ships for the engines on this path (read from the reference tree
    when it is there; the tree does not travel to the GPU box)."""
    import glob
    import os
    root = "/root/reference/config"
    paths = sorted(glob.glob(root + "/nearest_neighbor/*.json") + glob.glob(root + "/recommender/nearest_neighbor_recommender*.json") +
                   glob.glob(root + "/anomaly/light_lof*.json") + [root + "/classifier/nn.json"])
    paths = [p for p in paths if os.path.exists(p)]
    if not paths:
        pytest.skip("reference tree not present")
    d = nn.Datum().add_string("name", "juba").add_number("age", 31).add_number("w", 0.5)
    for p in paths:
        cfg = json.load(open(p))
        sfv = nn.convert(json.dumps(cfg["converter"]), d)
        keys = [k for k, _ in sfv]
        assert "age@num" in keys, p
        # string features: "str" tokens everywhere but default.json (bigram/idf: weight 0 on an empty model)
        assert "name$juba@str#bin/bin" in keys or p.endswith("nearest_neighbor/default.json"), p


def test_string_types_sample_weights_and_idf():
    """config/nearest_neighbor/default.json: bigram tokens, tf sample weight, idf global weight.
    Stored rows move the document counts, queries only read them:
    idf = log((documents + 1) / (documents with the feature + 1))."""
    import math
    conv = {"string_types": {"bigram": {"method": "ngram", "char_num": "2"}},
            "string_rules": [{"key": "*", "type": "bigram", "sample_weight": "tf", "global_weight": "idf"}],
            "num_types": {}, "num_rules": []}
    got = nn.convert_weighted(json.dumps(conv), [("t", "abab"), ("t", "abc"), ("t", "xyz")], ("t", "abcab"))
    want = [("t$ab@bigram#tf/idf", 2 * math.log(4 / 3)), ("t$bc@bigram#tf/idf", math.log(4 / 2)), ("t$ca@bigram#tf/idf", math.log(4 / 1))]
    assert [k for k, _ in got] == [k for k, _ in want]            # tokens of a value in sorted order
    for (_, g), (_, x) in zip(got, want):
        assert abs(g - x) < 1e-6
    # a feature every stored document has weighs log(1) = 0 and is dropped
    got = nn.convert_weighted(json.dumps(conv), [("t", "ab"), ("t", "ab")], ("t", "abc"))
    assert [k for k, _ in got] == ["t$bc@bigram#tf/idf"]
    # bin / log_tf sample weights, the space splitter, multi-byte characters
    conv2 = {"string_types": {"uni": {"method": "ngram", "char_num": 1}},
             "string_rules": [{"key": "w*", "type": "space", "sample_weight": "log_tf", "global_weight": "bin"},
                              {"key": "u*", "type": "uni", "sample_weight": "bin", "global_weight": "bin"}],
             "num_rules": []}
    got = nn.convert_weighted(json.dumps(conv2), [], ("words", "b a  b\tb"))
    assert got[0][0] == "words$a@space#log_tf/bin" and abs(got[0][1] - math.log(2)) < 1e-6
    assert got[1][0] == "words$b@space#log_tf/bin" and abs(got[1][1] - math.log(4)) < 1e-6
    got = nn.convert_weighted(json.dumps(conv2), [], ("u", "aあa"))
    assert got == [("u$a@uni#bin/bin", 1.0), ("u$あ@uni#bin/bin", 1.0)]
    for bad in ({"string_types": {"m": {"method": "mecab"}}}, {"string_rules": [{"key": "*", "type": "nope", "sample_weight": "bin", "global_weight": "bin"}]},
                {"string_rules": [{"key": "*", "type": "str", "sample_weight": "bin", "global_weight": "weight"}]}):
        with pytest.raises(nn.JubatusError):
            nn.convert_weighted(json.dumps(dict(conv, **bad)), [], ("t", "x"))
