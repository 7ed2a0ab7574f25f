"""Two answerers in parallel, then a refiner: the reference's three-model pipeline on the stage path.

The reference (Code/Quantised Models/combiner_quant_updated.py:177-203) loads Phi-2, Pythia-1B and
Llama-3.2-1B-Instruct in 8 bit, asks both answerers the question ONE AFTER THE OTHER on one GPU
(`get_response`, :89-113: sampling with temperature / top_k / top_p / repetition_penalty, answer = the decoded
sequence, tokens/s = new tokens / wall time), and lets the third model merge the two answers (`refine_summary`,
:116-138: temperature 0.5, top_k 30, top_p 0.9, repetition_penalty 1.1).  Only strings cross from model to
model.  The two answers do not depend on each other, so here each answerer is a `Pipeline` on its own GPU(s)
and the two generations run at the same time: one host thread advances both `Pipeline.generate_steps`
generators in turn, each step only enqueues work on that pipeline's GPU, so both GPUs stay busy (every `Stage`
entry point runs on its own device whatever device is current); the refiner follows on a third GPU or on one
of the two.

Tokenizers are the caller's (Hugging Face `AutoTokenizer` in the reference; any object with `encode(str) ->
ids` and `decode(ids) -> str`).  The prompt wording is a constructor argument: pass the reference's own
templates (:90-94, :117-125) to reproduce its prompts exactly.
"""
from __future__ import annotations

import time

import torch

ANSWER_PROMPT = ("You are a helpful assistant. Answer the question below in detail, in at least 50 words, with the relevant "
                 "facts and the terms a reader would expect.\n\nQuestion: {question}\nAnswer:")
REFINE_PROMPT = ("You are an expert assistant. Merge the best information of the two responses below into one clear, informative "
                 "answer of at least 50 words; avoid vague phrases and keep factual terms and named entities.\n\n"
                 "Response 1:\n{ans1}\n\nResponse 2:\n{ans2}\n\nReference (optional):\n{reference}\n\nFinal refined response:")


class Ensemble:
    """answerers: two (or more) `(pipeline, tokenizer)` pairs; refiner: one `(pipeline, tokenizer)` pair."""

    def __init__(self, answerers, refiner, answer_prompt: str = ANSWER_PROMPT, refine_prompt: str = REFINE_PROMPT):
        if len(answerers) < 2:
            raise ValueError("the ensemble needs at least two answerers")
        self.answerers, self.refiner = list(answerers), refiner
        self.answer_prompt, self.refine_prompt = answer_prompt, refine_prompt

    @staticmethod
    def _generate(pipe, tok, prompt: str, max_new_tokens: int, **sampling):
        """One `get_response` (combiner_quant_updated.py:89-113): (decoded text, new tokens per second)."""
        return Ensemble._generate_many([(pipe, tok)], prompt, max_new_tokens, [sampling])[0]

    @staticmethod
    def _generate_many(pairs, prompt: str, max_new_tokens: int, samplings):
        """`get_response` for several (pipeline, tokenizer) pairs AT THE SAME TIME from one host thread: every
        pipeline's `generate_steps` only enqueues work on its own GPU(s), so advancing the generators in turn keeps all
        of them busy; a pipeline's tokens/s is its new tokens over the time from the common start to ITS last token."""
        its, ids_len, out = [], [], [None] * len(pairs)
        for pipe, tok in pairs:
            torch.cuda.synchronize(pipe.stages[0].device)
        t0 = time.perf_counter()
        for (pipe, tok), smp in zip(pairs, samplings):
            ids = torch.as_tensor(tok.encode(prompt), dtype=torch.int64)
            ids_len.append(int(ids.shape[0]))
            its.append(pipe.generate_steps(ids, max_new_tokens=max_new_tokens, do_sample=True,
                                           eos_token_id=getattr(tok, "eos_token_id", None), **smp))
        live = list(range(len(pairs)))
        while live:
            for k in list(live):
                try:
                    next(its[k])
                except StopIteration as done:
                    seq = done.value.cpu()                      # waits for this pipeline's GPU only
                    dt = time.perf_counter() - t0
                    n_new = int(seq.shape[0]) - ids_len[k]
                    out[k] = (pairs[k][1].decode(seq.tolist()).strip(), n_new / dt if dt > 0 else 0.0)
                    live.remove(k)
        return out

    def answers(self, question: str, max_new_tokens: int = 200, temperature: float = 0.7, top_k: int = 50, top_p: float = 0.9,
                repetition_penalty: float = 1.1, seed: int = 0, concurrent: bool = True):
        """All answerers on the same question; returns [(answer, tokens_per_second)] in answerer order.
        concurrent=False runs them one after the other, as the reference does (:197, :200)."""
        prompt = self.answer_prompt.format(question=question.strip())
        smp = [dict(temperature=temperature, top_k=top_k, top_p=top_p, repetition_penalty=repetition_penalty, seed=seed + i)
               for i in range(len(self.answerers))]
        if concurrent:
            return self._generate_many(self.answerers, prompt, max_new_tokens, smp)
        return [self._generate_many([pr], prompt, max_new_tokens, [sm])[0] for pr, sm in zip(self.answerers, smp)]

    def refine(self, ans1: str, ans2: str, reference: str | None, max_new_tokens: int = 200, seed: int = 0) -> str:
        """`refine_summary` (combiner_quant_updated.py:116-138) with its fixed sampling parameters."""
        pipe, tok = self.refiner
        prompt = self.refine_prompt.format(ans1=ans1, ans2=ans2, reference=reference if reference else "N/A")
        text, _ = self._generate(pipe, tok, prompt, max_new_tokens, temperature=0.5, top_k=30, top_p=0.9, repetition_penalty=1.1,
                                 seed=seed)
        return text

    def __call__(self, question: str, reference: str | None = None, max_new_tokens: int = 200, **sampling) -> dict:
        """One iteration of the reference's evaluation loop (:191-212) up to the refined answer: the answers, the
        refined text and `tps` = the mean of the answerers' tokens/s (the refiner is not counted there either)."""
        seed = int(sampling.pop("seed", 0))
        concurrent = bool(sampling.pop("concurrent", True))
        res = self.answers(question, max_new_tokens, seed=seed, concurrent=concurrent, **sampling)
        refined = self.refine(res[0][0], res[1][0], reference, max_new_tokens, seed=seed)
        return {"answers": [a for a, _ in res], "refined": refined, "tps": sum(t for _, t in res) / len(res),
                "tps_each": [t for _, t in res]}
