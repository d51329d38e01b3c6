"""Result files and log markers of the reference's decoding stage, for the n-best lists this package returns.

Mirrors the tail of ``Speech2Text._decode_single_sample`` (``espnet2/bin/asr_inference.py:653-676``: drop
sos/eos, drop blank id 0, ids -> tokens -> text) and the writer loop of ``inference`` (``:895-907``:
``<outdir>/<n>best_recog/{token,token_int,score,text}``, one ``key value`` line per utterance, the layout of
``DatadirWriter``), plus the two log lines ``utils/calculate_rtf.py:68-86`` parses ("speech length: <samples>"
at the start and "best hypo: ..." at the end of an utterance), so ``asr.sh`` stages 12-13 can score the output
unchanged.  Host-side only; nothing here touches the device.
"""

import logging
import os
from typing import Callable, Dict, List, Optional, Sequence, Tuple


def hypothesis_to_result(hyp, token_list: Optional[Sequence[str]] = None,
                         tokens2text: Optional[Callable[[List[str]], str]] = None,
                         use_transducer_decoder: bool = False) -> Tuple[Optional[str], List[str], List[int], object]:
    """``(text, token, token_int, hyp)`` as ``Speech2Text.__call__`` returns it (``asr_inference.py:653-676``)."""
    yseq = hyp.yseq.tolist() if hasattr(hyp.yseq, "tolist") else list(hyp.yseq)
    token_int = yseq[1:] if use_transducer_decoder else yseq[1:-1]  # remove sos/eos
    token_int = [int(t) for t in token_int if t != 0]  # blank symbol id is assumed to be 0
    token = [token_list[t] for t in token_int] if token_list is not None else [str(t) for t in token_int]
    text = tokens2text(token) if tokens2text is not None else None
    return text, token, token_int, hyp


class NBestWriter:
    """``with NBestWriter(outdir) as w: w.write(key, nbest)`` -> ``<outdir>/<n>best_recog/<name>`` files."""

    def __init__(self, output_dir: str, token_list: Optional[Sequence[str]] = None,
                 tokens2text: Optional[Callable[[List[str]], str]] = None, nbest: int = 1):
        self.dir = output_dir
        self.token_list = token_list
        self.tokens2text = tokens2text
        self.nbest = nbest
        self._files: Dict[Tuple[int, str], object] = {}
        self._keys = set()

    def _file(self, n: int, name: str):
        f = self._files.get((n, name))
        if f is None:
            d = os.path.join(self.dir, f"{n}best_recog")
            os.makedirs(d, exist_ok=True)
            f = self._files[(n, name)] = open(os.path.join(d, name), "w", encoding="utf-8")
        return f

    def write(self, key: str, hyps) -> List[Tuple[Optional[str], List[str], List[int], object]]:
        if key in self._keys:  # DatadirWriter refuses duplicated keys
            raise RuntimeError(f"Duplicated: {key}")
        self._keys.add(key)
        results = [hypothesis_to_result(h, self.token_list, self.tokens2text) for h in hyps[: self.nbest]]
        for n, (text, token, token_int, hyp) in enumerate(results, 1):
            self._file(n, "token").write(f"{key} {' '.join(token)}\n")
            self._file(n, "token_int").write(f"{key} {' '.join(map(str, token_int))}\n")
            self._file(n, "score").write(f"{key} {float(hyp.score)}\n")
            if text is not None:
                self._file(n, "text").write(f"{key} {text}\n")
        return results

    def close(self):
        for f in self._files.values():
            f.close()
        self._files = {}

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def log_rtf_markers(speech_samples: int, best_text: str, logger: Optional[logging.Logger] = None) -> None:
    """The two INFO lines ``utils/calculate_rtf.py`` pairs up per utterance (espnet2 markers)."""
    lg = logger or logging.getLogger()
    lg.info("speech length: " + str(int(speech_samples)))
    lg.info("best hypo: " + best_text + "\n")
