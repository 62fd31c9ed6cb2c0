"""The GBNF engine on the grammar files that ship with the reference (`llama.cpp/grammars/*.gbnf`): every file parses, and sentences that
are / are not in each language are told apart.  Skipped where the reference tree is not present."""
import os

import pytest

from distributed_llm_pipeline_b200.serving.grammar import Grammar

DIR = os.environ.get("DLP_REFERENCE_GRAMMARS", "/root/reference/llama.cpp/grammars")
CASES = {
    # operands carry their trailing white space; an operator is followed by its operand at once
    "arithmetic": [("1 +2 *3 = 7\n", True), ("x = (1 +y) *3\n", False), ("(1 +y) *3 = x\n", True), ("1 ++2 = 3\n", False), ("1 +2 = 3", False)],
    "json": [('{"a": [1, 2.5, true, null], "b": {"c": "d\\n"}}', True), ('{"a": }', False), ("[1,2", False), ('"x"', False), ("{}", True)],
    "json_arr": [('[\n{"a": 1},\n2]', True), ('{"a": 1}', False), ("[1]", False)],
    "list": [("- one\n- two\n", True), ("one\n", False), ("- one", False)],
    "chess": [("1. e4 e5\n2. Nf3 Nc6\n", True), ("1. e4 e5\n", False), ("1. zz e5\n2. Nf3 Nc6\n", False), ("1. O-O O-O-O\n2. exd5 Qxd5+\n", True)],
    "japanese": [("こんにちは。", True), ("hello", False)],
    "c": [("int main() { return 0 }\n", False), ("", True)],      # root is a possibly empty list of declarations
    "english": [("Hello world.", True)],
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_reference_grammar_file(name):
    path = os.path.join(DIR, name + ".gbnf")
    if not os.path.exists(path):
        pytest.skip("reference grammar files are not present")
    g = Grammar(open(path, encoding="utf-8").read())
    for text, ok in CASES[name]:
        assert g.matches(text) == ok, (name, text)
