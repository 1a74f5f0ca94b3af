"""Generate tests/golden/spm_vocab.json: a small SentencePiece BPE vocabulary with byte fallback (the kind of
vocabulary "llama" GGUF files carry) plus encodings/decodings produced by the sentencepiece library itself.

The tokenizer in aios_b200/csrc/tokenizer.cpp is then checked against these ids (tests/test_cpu_runtime.py).
sentencepiece is llama.cpp's own ground truth for tokenizer.ggml.model == "llama"; the reference (aiOS) has no
tokenizer of its own — llama-server does it.

Run:  python scripts/make_golden_spm.py       (needs the sentencepiece wheel; deterministic)
"""
import io
import json
import os

import sentencepiece as spm

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "tests", "golden", "spm_vocab.json")

CORPUS = """
The agent orchestrator decomposes a goal into tasks and routes each task to an agent.
Each agent asks the runtime for inference at an intelligence level: reactive, operational, tactical or strategic.
The runtime loads a model, keeps it resident, and answers with text, tokens used and latency in milliseconds.
A model manager tracks the status of every model: loading, ready, error or unloading.
Health checks report how many models are loaded and the uptime of the service in seconds.
System prompts describe the role of the assistant; user prompts carry the question.
{"action": "restart_service", "service": "nginx", "reason": "memory usage above threshold"}
{"tool": "fs.read", "args": {"path": "/etc/hosts"}, "confidence": 0.93}
Please respond with a JSON object containing the fields thought, tool and args.
Network latency was 12.5 ms at 09:41; CPU load 0.75, 1.20, 3.05; disk /dev/sda1 is 87% full.
def load_model(name, path, context_length=2048): return manager.load(name, path, context_length)
for i in range(10): total += values[i] * weights[i]
What is the capital of France? The capital of France is Paris.
Hello, world! How are you today? I am fine, thank you very much.
""".strip().splitlines()

SAMPLES = [
    "Hello, world!",
    "The runtime loads a model and answers with text.",
    " leading space and  double  spaces ",
    "",
    " ",
    "tabs\tand\nnewlines\n\nhere",
    '{"tool": "fs.read", "args": {"path": "/tmp/x"}}',
    "unseen çharacters: naïve café — 你好 🙂",
    "for i in range(3): print(i)",
    "UPPER lower MiXeD 12345 67.89%",
    "aaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaa",
    "What is the capital of France?",
]


def main():
    model = io.BytesIO()
    spm.SentencePieceTrainer.train(
        sentence_iterator=iter(CORPUS * 4), model_writer=model, vocab_size=600, model_type="bpe", byte_fallback=True,
        character_coverage=1.0, unk_id=0, bos_id=1, eos_id=2, pad_id=-1, add_dummy_prefix=True,
        remove_extra_whitespaces=False, normalization_rule_name="identity", split_digits=True,
        allow_whitespace_only_pieces=True, num_threads=1)
    sp = spm.SentencePieceProcessor(model_proto=model.getvalue())
    vocab = []
    for i in range(sp.get_piece_size()):
        if sp.is_unknown(i):
            t = 2
        elif sp.is_control(i):
            t = 3
        elif sp.is_byte(i):
            t = 6
        elif sp.is_unused(i):
            t = 5
        else:
            t = 1
        vocab.append([sp.id_to_piece(i), sp.get_score(i), t])
    cases = []
    for s in SAMPLES:
        ids = sp.encode(s)
        cases.append({"text": s, "ids": ids, "decoded": sp.decode(ids)})
    with open(OUT, "w", encoding="utf-8") as f:
        json.dump({"generator": "scripts/make_golden_spm.py", "sentencepiece": spm.__version__, "vocab": vocab,
                   "cases": cases}, f, ensure_ascii=False, indent=0)
    print(f"wrote {OUT}: {len(vocab)} pieces, {len(cases)} cases")


if __name__ == "__main__":
    main()
