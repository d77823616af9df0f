"""Golden vectors for the host-side request logic of `infer_batch_process` (infer/utils_infer.py:503-770), produced by
the reference's OWN statements: the module imports packages that are absent here (torchaudio codecs, pydub, vocos,
transformers pipelines ...), so the three blocks are lifted out of the function body with `ast` and executed unmodified
in a namespace that supplies their free variables.

  * tokenizer dispatch (process_batch, utils_infer.py:549-676): which `convert_char_to_*` the reference calls, with
    which keyword arguments, and whether `normalize_n2gk_plus` runs first, for every tokenizer type / vocab flavour /
    skipTC / N2gk+ combination.  The converters are replaced by recording stubs: the RULE is pinned, the converters
    themselves stay the reference's Python.
  * duration rule (utils_infer.py:541-545, 678-685): local_speed / utf-8 byte lengths / fix_duration.
  * cross-fade join (utils_infer.py:734-770).

Run in the build container (needs /root/reference):  python oracle/make_golden_request.py
  -> tests/golden/request_level.json
"""
import ast
import itertools
import json
import os
import sys
import types

import numpy as np

REF = "/root/reference/src/f5_tts/infer/utils_infer.py"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _find(node, pred):
    for n in ast.walk(node):
        if pred(n):
            return n
    raise LookupError("block not found in the reference")


def lift_blocks():
    tree = ast.parse(open(REF, encoding="utf-8").read())
    outer = _find(tree, lambda n: isinstance(n, ast.FunctionDef) and n.name == "infer_batch_process")
    pb = _find(outer, lambda n: isinstance(n, ast.FunctionDef) and n.name == "process_batch")
    body = pb.body
    # statements of process_batch up to (not including) `with torch.inference_mode():`
    cut = next(i for i, s in enumerate(body) if isinstance(s, ast.With))
    head = body[:cut]
    # the cross-fade block: `if cross_fade_duration <= 0: ... else: ...` inside `if generated_waves:`
    gw = _find(outer, lambda n: isinstance(n, ast.If) and isinstance(n.test, ast.Name) and n.test.id == "generated_waves")
    xf = next(s for s in gw.body if isinstance(s, ast.If))

    def comp(stmts):
        return compile(ast.fix_missing_locations(ast.Module(body=list(stmts), type_ignores=[])), REF, "exec")

    return comp(head), comp([xf])


class _Audio:
    def __init__(self, n):
        self.shape = (1, n)


def run_head(code, ref_text, gen_text, n_audio, speed, fix_duration, tokenizer_type, vocab, skip_tc, n2gk, legacy):
    calls = []

    def stub(name):
        def f(text_list, **kw):
            calls.append({"fn": name, "kwargs": {k: (v if not isinstance(v, (set, frozenset, list, tuple)) else str(k))
                                                  for k, v in kw.items()}, "n": len(text_list)})
            return [list(t) for t in text_list]
        return f

    def n2gk_stub(t):
        calls.append({"fn": "normalize_n2gk_plus"})
        return t

    # sentinel filter constants: the dispatch passes them by name
    mu = types.ModuleType("f5_tts.model.utils")
    for nm in ("PHONEMES_N", "PHONEMES_I_NO_H", "PHONEMES_C_SONORANT"):
        setattr(mu, nm, nm)
    conv_names = ["convert_char_to_grapheme", "convert_char_to_grapheme_skipTC", "convert_char_to_allophone",
                  "convert_char_to_allophone_skipTC", "convert_char_to_phoneme", "convert_char_to_phoneme_skipTC",
                  "convert_char_to_pinyin", "convert_char_to_no_ieung_g2p", "convert_char_to_no_ieung_raw"]
    for nm in conv_names:
        setattr(mu, nm, stub(nm))
    saved = {k: sys.modules.get(k) for k in ("f5_tts", "f5_tts.model", "f5_tts.model.utils")}
    sys.modules["f5_tts"] = types.ModuleType("f5_tts")
    sys.modules["f5_tts.model"] = types.ModuleType("f5_tts.model")
    sys.modules["f5_tts.model.utils"] = mu
    model = types.SimpleNamespace(vocab_char_map=vocab, _use_skip_tc=skip_tc, _use_n2gk_plus=n2gk,
                                  _tokenizer_type=tokenizer_type, _tokenizer_version_legacy=legacy)
    ns = {nm: getattr(mu, nm) for nm in conv_names}
    ns.update(normalize_n2gk_plus=n2gk_stub, model_obj=model, ref_text=ref_text, gen_text=gen_text, speed=speed,
              fix_duration=fix_duration, audio=_Audio(n_audio), hop_length=256, target_sample_rate=24000,
              print=lambda *a, **k: None)
    try:
        exec(code, ns)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    # filters were passed as sentinel strings: keep their names
    for c in calls:
        for k, v in list(c.get("kwargs", {}).items()):
            if isinstance(v, str) and v in ("coda_filter", "initial_filter"):
                c["kwargs"][k] = v
    return calls, int(ns["duration"]), int(ns["ref_audio_len"])


VOCABS = {
    "allophone": {" ": 0, "k": 1, "kⁱ": 2, "pᶜ": 3, "a": 4},
    "phoneme": {" ": 0, "ㄱ": 1, "ㅏ": 2},
    "grapheme": {" ": 0, "ㅄ": 1, "ㅏ": 2, "ㄱ": 3},  # also contains ㄱ: phoneme rule fires first for "custom"
    "grapheme_only": {" ": 0, "ㅄ": 1, "ㅏ": 2},
    "latin": {" ": 0, "a": 1, "b": 2},
}
TYPES = ["custom", "kor_grapheme", "kor_allophone", "kor_phoneme", "kor_i_only", "kor_c_only", "kor_i_and_c",
         "kor_n_only", "kor_i_and_n", "kor_efficient_allophone", "kor_inf", "kor_nf", "kor_no_ieung_g2p",
         "kor_no_ieung_raw", "pinyin"]


def main():
    head, xf = lift_blocks()
    out = {"dispatch": [], "duration": [], "cross_fade": []}
    for ttype, (vname, vocab), skip, n2gk in itertools.product(TYPES, list(VOCABS.items()) + [("none", None)],
                                                               (False, True), (False, True)):
        calls, _, _ = run_head(head, "ref text. ", "generate this.", 72000, 1.0, None, ttype, vocab, skip, n2gk, False)
        # the filter sentinels come back as their own names; sets never occur
        for c in calls:
            c.pop("n", None)
        out["dispatch"].append({"tokenizer_type": ttype, "vocab": vname, "skip_tc": skip, "n2gk": n2gk, "calls": calls})
    # legacy flag only changes the skipTC converters' kwarg
    calls, _, _ = run_head(head, "r. ", "g", 72000, 1.0, None, "kor_allophone", VOCABS["allophone"], True, True, True)
    for c in calls:
        c.pop("n", None)
    out["dispatch"].append({"tokenizer_type": "kor_allophone", "vocab": "allophone", "skip_tc": True, "n2gk": True,
                            "legacy": True, "calls": calls})
    for ref_text, gen_text, n_audio, speed, fix in [
            ("some call me nature. ", "hello there, how are you today?", 72192, 1.0, None),
            ("안녕하세요. ", "오늘은 날씨가 정말 좋네요, 산책을 가고 싶어요!", 80000, 1.0, None),
            ("ref. ", "short", 72000, 1.0, None),          # < 10 bytes -> local_speed 0.3
            ("ref. ", "한국어", 72000, 1.0, None),           # 9 bytes utf-8 -> 0.3
            ("ref. ", "한국어!", 72000, 1.0, None),          # 10 bytes -> speed
            ("a longer reference text. ", "a generated sentence of some length.", 100001, 0.7, None),
            ("a longer reference text. ", "a generated sentence of some length.", 100001, 1.3, None),
            ("ref. ", "whatever text", 72000, 1.0, 12.5),
            ("ref. ", "x", 50000, 2.0, 3.0)]:
        _, dur, ral = run_head(head, ref_text, gen_text, n_audio, speed, fix, "pinyin", None, False, False, False)
        out["duration"].append({"ref_text": ref_text, "gen_text": gen_text, "n_audio": n_audio, "speed": speed,
                                "fix_duration": fix, "ref_audio_len": ral, "duration": dur})
    rng = np.random.default_rng(0)
    # (small rates keep the fixture small: the block only uses int(cross_fade_duration * target_sample_rate))
    for lens, dur, sr in [((500, 700, 600), 0.15, 1000), ((500, 700), 0.0, 1000), ((300, 100, 500), 0.15, 1000),
                          ((400,), 0.15, 1000), ((200, 200, 200, 200), 0.05, 1000), ((100, 0, 50), 0.15, 1000),
                          ((300, 300), -1.0, 1000), ((4000, 3700), 0.15, 24000)]:
        waves = [rng.standard_normal(n).astype(np.float32) for n in lens]
        ns = {"np": np, "generated_waves": [w.copy() for w in waves], "cross_fade_duration": dur,
              "target_sample_rate": sr}
        exec(xf, ns)
        out["cross_fade"].append({"lens": list(lens), "cross_fade_duration": dur, "sample_rate": sr,
                                  "seed_note": "default_rng(0) stream, float32 draws in this order", "final": np.asarray(ns["final_wave"], dtype=np.float64).tolist()})
    path = os.path.join(ROOT, "tests", "golden", "request_level.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=False)
    print(path, os.path.getsize(path), "bytes;", len(out["dispatch"]), "dispatch rows")


if __name__ == "__main__":
    main()
