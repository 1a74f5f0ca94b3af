"""Generate tests/golden/chat_templates.json: chat templates of the kinds GGUF files carry in tokenizer.chat_template,
each rendered by the jinja2 package the way transformers (and llama.cpp's Jinja engine after it) renders them —
ImmutableSandboxedEnvironment(trim_blocks=True, lstrip_blocks=True), raise_exception available — for several message
lists.  aios_b200/csrc/jinja_lite.cpp must reproduce every rendering (or fail where jinja2 raises).

The templates are restated from the public model cards' tokenizer configurations (TinyLlama-1.1B-Chat = Zephyr format,
Mistral-7B-Instruct-v0.2, Meta-Llama-3-Instruct, ChatML, Llama-2-chat) plus two synthetic ones that exercise the rest of
the supported subset.  Run:  python scripts/make_golden_chat.py
"""
import json
import os

from jinja2.exceptions import TemplateError
from jinja2.sandbox import ImmutableSandboxedEnvironment

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "tests", "golden", "chat_templates.json")

TEMPLATES = {
    "zephyr": (
        "{% for message in messages %}\n{% if message['role'] == 'user' %}\n{{ '<|user|>\\n' + message['content'] + eos_token }}\n"
        "{% elif message['role'] == 'system' %}\n{{ '<|system|>\\n' + message['content'] + eos_token }}\n"
        "{% elif message['role'] == 'assistant' %}\n{{ '<|assistant|>\\n'  + message['content'] + eos_token }}\n{% endif %}\n"
        "{% if loop.last and add_generation_prompt %}\n{{ '<|assistant|>' }}\n{% endif %}\n{% endfor %}"),
    "mistral-v0.2": (
        "{{ bos_token }}{% for message in messages %}{% if (message['role'] == 'user') != (loop.index0 % 2 == 0) %}"
        "{{ raise_exception('Conversation roles must alternate user/assistant/user/assistant/...') }}{% endif %}"
        "{% if message['role'] == 'user' %}{{ '[INST] ' + message['content'] + ' [/INST]' }}"
        "{% elif message['role'] == 'assistant' %}{{ message['content'] + eos_token}}"
        "{% else %}{{ raise_exception('Only user and assistant roles are supported!') }}{% endif %}{% endfor %}"),
    "llama3": (
        "{% set loop_messages = messages %}{% for message in loop_messages %}"
        "{% set content = '<|start_header_id|>' + message['role'] + '<|end_header_id|>\\n\\n'+ message['content'] | trim + '<|eot_id|>' %}"
        "{% if loop.index0 == 0 %}{% set content = bos_token + content %}{% endif %}{{ content }}{% endfor %}"
        "{% if add_generation_prompt %}{{ '<|start_header_id|>assistant<|end_header_id|>\\n\\n' }}{% endif %}"),
    "chatml": (
        "{% for message in messages %}{{'<|im_start|>' + message['role'] + '\\n' + message['content'] + '<|im_end|>' + '\\n'}}"
        "{% endfor %}{% if add_generation_prompt %}{{ '<|im_start|>assistant\\n' }}{% endif %}"),
    "llama2": (
        "{% if messages[0]['role'] == 'system' %}{% set loop_messages = messages[1:] %}"
        "{% set system_message = messages[0]['content'] %}{% else %}{% set loop_messages = messages %}"
        "{% set system_message = false %}{% endif %}{% for message in loop_messages %}"
        "{% if (message['role'] == 'user') != (loop.index0 % 2 == 0) %}"
        "{{ raise_exception('Conversation roles must alternate user/assistant/user/assistant/...') }}{% endif %}"
        "{% if loop.index0 == 0 and system_message != false %}"
        "{% set content = '<<SYS>>\\n' + system_message + '\\n<</SYS>>\\n\\n' + message['content'] %}"
        "{% else %}{% set content = message['content'] %}{% endif %}"
        "{% if message['role'] == 'user' %}{{ bos_token + '[INST] ' + content.strip() + ' [/INST]' }}"
        "{% elif message['role'] == 'assistant' %}{{ ' '  + content.strip() + ' ' + eos_token }}{% endif %}{% endfor %}"),
    "synthetic-namespace": (
        "{%- set ns = namespace(n_user=0, last='') -%}\n"
        "{%- for m in messages -%}\n"
        "  {%- if m.role == 'user' -%}{%- set ns.n_user = ns.n_user + 1 -%}{%- endif -%}\n"
        "  {%- set ns.last = m.role -%}\n"
        "  [{{ loop.index }}/{{ loop.length }}:{{ m.role | upper }}] {{ m['content'] | trim }}{{ '' if loop.last else ' | ' }}\n"
        "{%- endfor %}\n"
        "users={{ ns.n_user }} last={{ ns.last }} n={{ messages | length }} first={{ (messages | first).role }}"
        "{% if messages | length > 1 and messages[-1].role != 'assistant' %} open{% endif %}\n"
        "{# a comment #}"
        "{{ 'gen' if add_generation_prompt else 'nogen' }} {{ 7 // 2 }} {{ -7 % 3 }} {{ 'ab' ~ 3 ~ none }} {{ 'x' in 'xyz' }} "
        "{{ undefined_name is defined }} {{ bos_token is string }} {{ messages[0].nothing | default('dflt') }}"),
    "synthetic-blocks": (
        "Header\n"
        "{% for m in messages %}\n"
        "    {% if loop.first %}\n"
        "first: {{ m.content }}\n"
        "    {% elif loop.last %}\n"
        "last: {{ m.content }}\n"
        "    {% else %}\n"
        "mid: {{ m.content[:5] }}...{{ m.content[-3:] }}\n"
        "    {% endif %}\n"
        "{% endfor %}\n"
        "{%+ if add_generation_prompt %}go{% endif %}\n"
        "Footer {{ messages | length * 2 + 1 }}\n"),
}

CONVERSATIONS = {
    "system+user": [("system", "Be brief."), ("user", "What is the capital of France?")],
    "user": [("user", "  Hello, world!  ")],
    "three-turns": [("user", "Hi"), ("assistant", "Hello! How can I help?"), ("user", "Tell me a joke.\nA short one.")],
    "system+three": [("system", "You are aiOS."), ("user", "status?"), ("assistant", "{\"ok\": true}"), ("user", "thanks")],
    "bad-order": [("assistant", "I speak first"), ("user", "no")],
}


def main():
    env = ImmutableSandboxedEnvironment(trim_blocks=True, lstrip_blocks=True)

    def raise_exception(msg):
        raise TemplateError(msg)

    env.globals["raise_exception"] = raise_exception
    cases = []
    for tname, src in TEMPLATES.items():
        tmpl = env.from_string(src)
        for cname, conv in CONVERSATIONS.items():
            msgs = [{"role": r, "content": c} for r, c in conv]
            for gen in (True, False):
                try:
                    text = tmpl.render(messages=msgs, bos_token="<s>", eos_token="</s>", add_generation_prompt=gen)
                except TemplateError as e:
                    text = None
                cases.append({"template": tname, "conversation": cname, "add_generation_prompt": gen, "expected": text})
    out = {"note": "generated by scripts/make_golden_chat.py with jinja2 " + __import__("jinja2").__version__,
           "templates": TEMPLATES, "conversations": CONVERSATIONS, "cases": cases}
    with open(OUT, "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=False, indent=0)
    n_err = sum(c["expected"] is None for c in cases)
    print(f"{OUT}: {len(TEMPLATES)} templates x {len(CONVERSATIONS)} conversations x 2 = {len(cases)} cases ({n_err} raise)")


if __name__ == "__main__":
    main()
