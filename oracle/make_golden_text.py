"""Golden vectors for the request-level text chunker: the reference's own `chunk_text` (infer/utils_infer.py:83-110) is
lifted out of its module with `ast` (the module itself imports packages absent here) and run on fixed inputs.
Run in the build container (needs /root/reference): python oracle/make_golden_text.py -> tests/golden/chunk_text.json"""
import ast
import json
import os
import re  # noqa: F401  (used by the lifted function)

REF = "/root/reference/src/f5_tts/infer/utils_infer.py"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = [
    ("Hello there. This is a test, with several clauses; some short: some long! Does it split? Yes.", 40),
    ("안녕하세요. 오늘은 날씨가 정말 좋네요, 산책을 가고 싶어요! 같이 갈래요? 좋아요.", 60),
    ("中文句子。没有空格的标点，也要切分！对吗？对。", 30),
    ("no punctuation at all just one very long run of words that exceeds the budget", 20),
    ("", 50),
    ("Short.", 135),
    ("A. B. C. D. E. F. G. H. I. J.", 6),
    ("Mixed 한국어 and English, 섞어서 씁니다. OK? 네！끝。", 25),
]


def main():
    tree = ast.parse(open(REF, encoding="utf-8").read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "chunk_text")
    ns = {"re": re}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), REF, "exec"), ns)
    out = [{"text": t, "max_chars": m, "chunks": ns["chunk_text"](t, max_chars=m)} for t, m in CASES]
    with open(os.path.join(ROOT, "tests", "golden", "chunk_text.json"), "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=False, indent=1)
    print(len(out), "cases")


if __name__ == "__main__":
    main()
