"""Which text converter a request uses -- the rule of `infer_batch_process.process_batch`
(reference src/f5_tts/infer/utils_infer.py:549-676), as a table.

The converters themselves (`convert_char_to_allophone`, `..._grapheme`, `..._phoneme`, `..._pinyin`, N2gk+
normalisation; reference src/f5_tts/model/utils.py:152-518) stay the reference's host-side Python, exactly as
`north_star` and SURVEY.md 8(b) prescribe: they are looked up in `f5_tts.model.utils` at call time.  What lives here is
only the DISPATCH -- pinned row by row to the reference's own statements by `tests/golden/request_level.json`
(`oracle/make_golden_request.py`, 361 combinations) -- so that a Korean allophone / jamo checkpoint loaded through
`load_model` can never silently receive per-character tokens: when the converters are not importable and the
caller gave no `tokenize=` callable, the request fails loudly."""
from __future__ import annotations

from typing import Callable, List, NamedTuple, Optional

ALLOPHONE_MARKS = ("ⁱ", "ᶜ", "ʲ")  # superscripts that only an allophone vocabulary contains (utils_infer.py:570)

# tokenizer types that always go through convert_char_to_allophone with fixed switches (utils_infer.py:582-636):
#   name -> (apply_init, apply_coda, coda_filter constant, initial_filter constant); apply_pal is always False
_ALLOPHONE_MODES = {
    "kor_i_only": (True, False, None, None),
    "kor_c_only": (False, True, None, None),
    "kor_i_and_c": (True, True, None, None),
    "kor_n_only": (False, True, "PHONEMES_N", None),
    "kor_i_and_n": (True, True, "PHONEMES_N", None),
    "kor_efficient_allophone": (True, True, "PHONEMES_C_SONORANT", "PHONEMES_I_NO_H"),
    "kor_inf": (True, True, "PHONEMES_C_SONORANT", None),
    "kor_nf": (False, True, "PHONEMES_C_SONORANT", None),
}
_FAMILY_FN = {"grapheme": "convert_char_to_grapheme", "allophone": "convert_char_to_allophone",
              "phoneme": "convert_char_to_phoneme"}


class Choice(NamedTuple):
    normalize: bool   # run normalize_n2gk_plus over every text first
    fn: str           # name of the converter in f5_tts.model.utils
    kwargs: dict      # keyword arguments; filter constants are given by NAME (resolved against the same module)


def _family(tokenizer_type: str, vocab: dict) -> Optional[str]:
    """grapheme / allophone / phoneme, by explicit type first, then by what the vocabulary contains."""
    if tokenizer_type == "kor_grapheme":
        return "grapheme"
    if tokenizer_type == "kor_allophone":
        return "allophone"
    custom = tokenizer_type == "custom"
    if custom and any(m in k for k in vocab for m in ALLOPHONE_MARKS):
        return "allophone"
    if tokenizer_type in _ALLOPHONE_MODES or tokenizer_type in ("kor_no_ieung_g2p", "kor_no_ieung_raw"):
        return None
    if tokenizer_type == "kor_phoneme" or (custom and "ㄱ" in vocab):
        return "phoneme"
    if "ㅄ" in vocab:
        return "grapheme"
    return None


def choose(tokenizer_type: str = "custom", vocab: Optional[dict] = None, use_skip_tc: bool = False,
           use_n2gk_plus: bool = True, legacy: bool = False) -> Choice:
    if vocab is None:  # no vocab map: pinyin, nothing else is consulted (utils_infer.py:673-675)
        return Choice(False, "convert_char_to_pinyin", {})
    fam = _family(tokenizer_type, vocab)
    if fam is not None:
        if use_skip_tc:
            return Choice(use_n2gk_plus, _FAMILY_FN[fam] + "_skipTC", {"legacy": legacy})
        return Choice(use_n2gk_plus, _FAMILY_FN[fam], {})
    if tokenizer_type in _ALLOPHONE_MODES:
        init, coda, coda_filter, initial_filter = _ALLOPHONE_MODES[tokenizer_type]
        return Choice(use_n2gk_plus, "convert_char_to_allophone",
                      dict(apply_init=init, apply_pal=False, apply_coda=coda, coda_filter=coda_filter,
                           initial_filter=initial_filter))
    if tokenizer_type in ("kor_no_ieung_g2p", "kor_no_ieung_raw"):
        return Choice(use_n2gk_plus, "convert_char_to_" + tokenizer_type[4:], {})
    return Choice(False, "convert_char_to_pinyin", {})


def choose_for_model(model_obj) -> Choice:
    """Read the attributes `load_model` stores on the model (utils_infer.py:339-343 -> :550-553)."""
    return choose(getattr(model_obj, "_tokenizer_type", "custom"), getattr(model_obj, "vocab_char_map", None),
                  getattr(model_obj, "_use_skip_tc", False), getattr(model_obj, "_use_n2gk_plus", True),
                  getattr(model_obj, "_tokenizer_version_legacy", False))


def char_tokens(text_list: List[str]) -> List[List[str]]:
    """One token per character, nothing else.  An explicit opt-in for plain character vocabularies: it is NOT what the
    reference's default `convert_char_to_pinyin` yields in general (that one re-spaces jieba segments and replaces CJK
    runs by pinyin syllables, model/utils.py:481-518)."""
    return [list(t) for t in text_list]


def _reference_utils():
    try:
        import f5_tts.model.utils as ru  # the reference's own module, when the drop-in runs inside its tree
        return ru
    except Exception:  # absent on this image (and its g2pk2 / pypinyin dependencies with it)
        return None


def resolve(choice: Choice, module=None) -> Callable[[List[str]], List[List[str]]]:
    """Bind a Choice to the reference's converter functions.  `module` defaults to `f5_tts.model.utils`."""
    ru = module if module is not None else _reference_utils()
    if ru is None:
        raise RuntimeError(f"this model's tokenizer is `{choice.fn}` (reference infer/utils_infer.py:549-676) but "
                           "f5_tts.model.utils is not importable: install the reference package next to f5b200 or "
                           "pass tokenize=<callable list[str] -> list[list[str]]> (f5b200.tokenizers.char_tokens for a plain "
                           "character vocabulary) -- refusing to fall back to characters silently")
    fn = getattr(ru, choice.fn)
    kwargs = {k: (getattr(ru, v) if isinstance(v, str) else v) for k, v in choice.kwargs.items()}
    norm = getattr(ru, "normalize_n2gk_plus") if choice.normalize else None

    def tokenize(text_list):
        if norm is not None:
            text_list = [norm(t) for t in text_list]
        return fn(text_list, **kwargs)
    return tokenize
