"""JSON-schema -> grammar against the reference's converter.  `llama-server` echoes the GBNF it built for a request's `json_schema`
(`generation_settings.grammar`); that text is (1) parsed and run by THIS repository's grammar engine - a parser test on real reference
output - and (2) compared, document by document, with the grammar `json_schema_to_gbnf` builds for the same schema: both must accept exactly
the documents that are valid.  (common/json-schema-to-grammar.cpp, llama-grammar.cpp)"""
import json
import os
import subprocess
import sys
import time
import urllib.request

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baseline import reference as ref_build  # noqa: E402

from distributed_llm_pipeline_b200.serving.grammar import Grammar, int_range_gbnf, json_schema_to_gbnf  # noqa: E402

PERSON = {"type": "object", "properties": {"name": {"type": "string", "minLength": 1, "maxLength": 5}, "age": {"type": "integer", "minimum": 0, "maximum": 120},
                                           "tags": {"type": "array", "items": {"type": "string"}, "minItems": 1, "maxItems": 3},
                                           "when": {"type": "string", "format": "date"}}, "required": ["name", "age"], "additionalProperties": False}
CASES = {
    "person": (PERSON, {'{"name":"bob","age":42}': True, '{"name":"bob","age":42,"tags":["a"]}': True, '{"name":"bob","age":42,"when":"2024-01-31"}': True,
                        '{"name":"bob","age":42,"tags":["a","b"],"when":"2024-12-01"}': True, '{"name":"","age":1}': False, '{"name":"toolong","age":1}': False,
                        '{"name":"bob","age":121}': False, '{"name":"bob"}': False, '{"name":"bob","age":3,"tags":[]}': False,
                        '{"name":"bob","age":3,"x":1}': False, '{"name":"bob","age":3,"when":"2024-13-01"}': False, '{"name":"bob","age":-1}': False}),
    "refs": ({"$defs": {"pt": {"type": "object", "properties": {"x": {"type": "integer"}, "y": {"type": "integer"}}, "required": ["x", "y"]}},
              "type": "object", "properties": {"a": {"$ref": "#/$defs/pt"}, "b": {"$ref": "#/$defs/pt"}}, "required": ["a"]},
             {'{"a":{"x":1,"y":-2}}': True, '{"a":{"x":1,"y":2},"b":{"x":0,"y":0}}': True, '{"a":{"x":1}}': False, '{"b":{"x":1,"y":2}}': False, '{"a":{"x":1.5,"y":2}}': False}),
    "anyof-enum": ({"anyOf": [{"type": "object", "properties": {"kind": {"const": "circle"}, "r": {"type": "number"}}, "required": ["kind", "r"]},
                              {"type": "object", "properties": {"kind": {"const": "square"}, "side": {"type": "number"}}, "required": ["kind", "side"]},
                              {"enum": ["none", 0, None]}]},
                   {'{"kind":"circle","r":1.5}': True, '{"kind":"square","side":2}': True, '"none"': True, '0': True, 'null': True,
                    '{"kind":"circle","side":2}': False, '{"kind":"triangle","r":1}': False, '"some"': False, '1': False}),
    "tuple-nullable": ({"type": "array", "prefixItems": [{"type": "string"}, {"type": ["integer", "null"]}, {"type": "boolean"}], "items": False, "minItems": 3, "maxItems": 3},
                       {'["a",1,true]': True, '["a",null,false]': True, '["a",1]': False, '[1,1,true]': False, '["a","b",true]': False}),
    "ranges": ({"type": "object", "properties": {"t": {"type": "integer", "minimum": -40, "maximum": 125}, "n": {"type": "integer", "exclusiveMinimum": 0},
                                                  "id": {"type": "string", "format": "uuid"}}, "required": ["t", "n", "id"]},
               {'{"t":-40,"n":1,"id":"123e4567-e89b-12d3-a456-426614174000"}': True, '{"t":125,"n":99999,"id":"00000000-0000-0000-0000-000000000000"}': True,
                '{"t":-41,"n":1,"id":"123e4567-e89b-12d3-a456-426614174000"}': False, '{"t":0,"n":0,"id":"123e4567-e89b-12d3-a456-426614174000"}': False,
                '{"t":0,"n":1,"id":"123e4567e89b12d3a456426614174000"}': False, '{"t":126,"n":1,"id":"123e4567-e89b-12d3-a456-426614174000"}': False}),
    "map": ({"type": "object", "additionalProperties": {"type": "integer"}},
            {'{}': True, '{"a":1}': True, '{"a":1,"b":-2}': True, '{"a":"x"}': False, '{"a":1.5}': False}),
    "allof": ({"allOf": [{"type": "object", "properties": {"a": {"type": "integer"}}, "required": ["a"]}, {"type": "object", "properties": {"b": {"type": "boolean"}}, "required": ["b"]}]},
              {'{"a":1,"b":true}': True, '{"a":1}': False, '{"b":true}': False, '{"a":"1","b":true}': False}),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_converter_accepts_exactly_the_valid_documents(name):
    schema, docs = CASES[name]
    g = Grammar(json_schema_to_gbnf(schema))
    for doc, ok in docs.items():
        json.loads(doc)
        assert g.matches(doc) == ok, (name, doc)
        if ok:       # white space where JSON allows it does not change the verdict
            spaced = json.dumps(json.loads(doc), indent=1)
            assert g.matches(spaced), (name, spaced)


def test_integer_ranges_exhaustively():
    for lo, hi in [(0, 120), (5, 5), (-17, 23), (None, -3), (100, None), (0, None), (None, 0), (995, 1005), (-1000, -10), (7, 123456)]:
        g = Grammar("root ::= " + int_range_gbnf(lo, hi) + '\ninteger ::= "-"? ( "0" | [1-9] [0-9]{0,15} )\n')
        a, b = (lo if lo is not None else -3000), (hi if hi is not None else 3000)
        for v in list(range(a - 150, min(b, a + 2500) + 150)) + list(range(max(a, b - 300), b + 150)) + [10 ** 6, -10 ** 6, 123456, 123457]:
            assert g.matches(str(v)) == ((lo is None or v >= lo) and (hi is None or v <= hi)), (lo, hi, v)
        assert not any(g.matches(x) for x in ["007", "+5", "1.0", ""])


@pytest.fixture(scope="module")
def ref_grammar(tmp_models, tmp_path_factory):
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-server"):
        pytest.skip(f"reference binaries unavailable: {why}")
    env = ref_build.cpu_env(str(tmp_path_factory.mktemp("refjs")))
    port = 38410
    p = subprocess.Popen([ref_build.exe("llama-server"), "-m", tmp_models[("tiny", "F16")], "-ngl", "0", "-c", "256", "--host", "127.0.0.1", "--port", str(port), "-t", "2",
                          "--no-warmup"], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    for _ in range(200):
        try:
            urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=1).read()
            break
        except Exception:
            time.sleep(0.1)
    else:
        p.terminate()
        pytest.skip("the reference server did not come up here")

    def get(schema):
        r = urllib.request.urlopen(urllib.request.Request(f"http://127.0.0.1:{port}/completion", json.dumps({"prompt": "x", "n_predict": 1, "json_schema": schema}).encode(),
                                                          {"Content-Type": "application/json"}), timeout=60)
        return json.loads(r.read())["generation_settings"]["grammar"]
    yield get
    p.terminate()
    p.wait()


@pytest.mark.parametrize("name", sorted(CASES))
def test_reference_grammar_runs_here_and_agrees(ref_grammar, name):
    schema, docs = CASES[name]
    try:
        text = ref_grammar(schema)
    except Exception as e:       # a schema feature the reference refuses is not a parity question
        pytest.skip(f"the reference did not convert this schema: {e}")
    theirs = Grammar(text)                 # the reference's GBNF, parsed and executed by this repository's engine
    ours = Grammar(json_schema_to_gbnf(schema))
    for doc, ok in docs.items():
        assert theirs.matches(doc) == ours.matches(doc) == ok, (name, doc, theirs.matches(doc), ours.matches(doc))
