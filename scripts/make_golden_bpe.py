"""Generate tests/golden/bpe_vocab.json: a small byte-level BPE vocabulary of the kind "gpt2" GGUF files carry
(Llama-3: tokenizer.ggml.model = "gpt2", tokenizer.ggml.pre = "llama-bpe") plus encodings produced by the HF
`tokenizers` library — the implementation Llama-3's own tokenizer.json runs on — with (a) the Llama-3 pre-tokenizer
pattern and ignore_merges, (b) GPT-2's pattern.

aios_b200/csrc/tokenizer.cpp is checked against these ids (tests/test_cpu_runtime.py).  The reference (aiOS) has no
tokenizer of its own: llama-server does it; real Llama-3 vocabularies cannot be fetched here, so the
vocabulary is trained on the spot from the corpus below (deterministic for a given tokenizers version; the fixture is
committed).

Run:  python scripts/make_golden_bpe.py
"""
import json
import os

from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "tests", "golden", "bpe_vocab.json")

LLAMA3_PAT = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+"
              r"|\s+(?!\S)|\s+")

CORPUS = """
The agent orchestrator decomposes a goal into tasks and routes each task to an agent.
Each agent asks the runtime for inference at an intelligence level: reactive, operational, tactical or strategic.
The runtime loads a model, keeps it resident, and answers with text, tokens used and latency in milliseconds.
A model manager tracks the status of every model: loading, ready, error or unloading.
Health checks report how many models are loaded and the uptime of the service in seconds.
It's the assistant's role; they're sure we've done it, I'm told you'll see, he'd agree. DON'T SHOUT, IT'S FINE.
{"action": "restart_service", "service": "nginx", "reason": "memory usage above threshold"}
{"tool": "fs.read", "args": {"path": "/etc/hosts"}, "confidence": 0.93}
Please respond with a JSON object containing the fields thought, tool and args.
Network latency was 12.5 ms at 09:41; CPU load 0.75, 1.20, 3.05; disk /dev/sda1 is 87% full; 1234567890 bytes.
def load_model(name, path, context_length=2048):
    return manager.load(name, path, context_length)

for i in range(10):
    total += values[i] * weights[i]
What is the capital of France? The capital of France is Paris.
Hello, world! How are you today? I am fine, thank you very much.
Der schnelle braune Fuchs springt über den faulen Hund. Größe, Maße, Straße.
Быстрая коричневая лиса перепрыгивает через ленивую собаку.
你好，世界！今天天气怎么样？我很好，谢谢。
naïve café déjà vu — “quoted” ‘text’ … 🙂🙂 👍
""".strip("\n").split("\n")

SAMPLES = [
    "Hello, world!",
    "The runtime loads a model and answers with text.",
    " leading space and  double  spaces ",
    "",
    " ",
    "  ",
    "tabs\tand\nnewlines\n\nhere",
    "trailing newline\n",
    "crlf line\r\nnext\r\n\r\nend",
    "blank   \n   lines \n\n\n  indented",
    '{"tool": "fs.read", "args": {"path": "/tmp/x"}}',
    "unseen çharacters: naïve café — 你好 🙂",
    "for i in range(3): print(i)",
    "UPPER lower MiXeD 12345 67.89% 1234567",
    "aaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaa",
    "What is the capital of France?",
    "It's we're I've I'm you'll he'd can't DON'T IT'S WE'LL 'tis 'Re",
    "x'sx 'd' ''s",
    "!!!\n\n???   ...\r\n",
    "a b c　d",          # no-break space, em space, ideographic space
    "Быстрая лиса 123 Größe Straße",
    "你好，世界！12345",
    "٣٤٥ ①②③ Ⅷ ½",                   # non-ASCII numbers (Nd, No, Nl)
    "emoji 🙂👍🏽 and é combining",
    "$$$ ### @@@user #tag <|x|>",
    " \n \n a",
    "end with spaces   ",
    "\n",
    "\n\n\n",
    " \t ",
]


def train(pre):
    tok = Tokenizer(models.BPE(ignore_merges=True))
    tok.pre_tokenizer = pre
    tok.decoder = decoders.ByteLevel()
    trainer = trainers.BpeTrainer(vocab_size=900, special_tokens=["<|begin_of_text|>", "<|end_of_text|>", "<|eot_id|>",
                                                                  "<|start_header_id|>", "<|end_header_id|>"],
                                  initial_alphabet=pre_tokenizers.ByteLevel.alphabet(), show_progress=False)
    tok.train_from_iterator(CORPUS * 4, trainer)
    return tok


def main():
    l3_pre = pre_tokenizers.Sequence([pre_tokenizers.Split(Regex(LLAMA3_PAT), behavior="isolated", invert=False),
                                      pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tok = train(l3_pre)
    spec = json.loads(tok.to_str())
    vocab = sorted(spec["model"]["vocab"].items(), key=lambda kv: kv[1])
    assert [i for _, i in vocab] == list(range(len(vocab)))
    merges = [m if isinstance(m, str) else " ".join(m) for m in spec["model"]["merges"]]
    cases_l3 = [[s, tok.encode(s, add_special_tokens=False).ids] for s in SAMPLES]
    for s, ids in cases_l3:
        assert tok.decode(ids, skip_special_tokens=False) == s, s
    # the same vocabulary and merges behind GPT-2's pattern, merges always applied
    gpt2 = Tokenizer(models.BPE(vocab=spec["model"]["vocab"], merges=[tuple(m.split(" ")) for m in merges]))
    gpt2.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=True)
    gpt2.decoder = decoders.ByteLevel()
    cases_gpt2 = [[s, gpt2.encode(s, add_special_tokens=False).ids] for s in SAMPLES]
    out = {"note": "generated by scripts/make_golden_bpe.py with tokenizers " + __import__("tokenizers").__version__,
           "vocab": [p for p, _ in vocab], "n_special": 5, "merges": merges,
           "cases_llama3": cases_l3, "cases_gpt2": cases_gpt2}
    with open(OUT, "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=False, indent=0)
    print(f"{OUT}: {len(vocab)} pieces, {len(merges)} merges, {len(SAMPLES)} cases x 2")


if __name__ == "__main__":
    main()
