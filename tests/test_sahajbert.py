"""sahajBERT pieces on CPU: Bengali tokenizer training + fast tokenizer class, lazy streaming dataset with SOP
instances, authorizer protocol (local authority + HF client with a fake hub), downstream fine-tuning loops."""
import asyncio
from datetime import timedelta

import pytest
import torch

BN_DOCS = [
    "আমি বাংলায় গান গাই। আমি বাংলার গান গাই। আমি আমার আমিকে চিরদিন এই বাংলায় খুঁজে পাই।",
    "ঢাকা বাংলাদেশের রাজধানী। এটি বুড়িগঙ্গা নদীর তীরে অবস্থিত। শহরটি ১৬১০ সালে প্রতিষ্ঠিত হয়।",
    "বিজ্ঞান হলো জ্ঞান অর্জনের একটি পদ্ধতি। পরীক্ষা ও পর্যবেক্ষণ এর ভিত্তি। বিজ্ঞানীরা প্রশ্ন করেন! উত্তর খোঁজেন?",
    "নদী পাহাড় থেকে নামে। সাগরে গিয়ে মেশে। পথে অনেক গ্রাম পড়ে। মানুষ নদীর জল ব্যবহার করে।",
] * 8


@pytest.fixture(scope="module")
def bn_tokenizer(tmp_path_factory):
    from dedloc_b200.trainers.sahajbert.tokenization_albert_bengali_fast import AlbertBengaliTokenizerFast
    from dedloc_b200.trainers.sahajbert.tokenizer.tokenizer_training_custom import SPECIAL_TOKENS, train_tokenizer

    out = tmp_path_factory.mktemp("tok") / "tokenizer.json"
    train_tokenizer(BN_DOCS, vocab_size=200, out_path=str(out), show_progress=False)
    tok = AlbertBengaliTokenizerFast(tokenizer_file=str(out))
    assert [tok.convert_tokens_to_ids(t) for t in SPECIAL_TOKENS] == [0, 1, 2, 3, 4]
    return tok


def test_tokenizer_class(bn_tokenizer):
    tok = bn_tokenizer
    assert tok.model_max_length == 512
    enc = tok("আমি বাংলায় গান গাই।", "ঢাকা রাজধানী।")
    assert enc["input_ids"][0] == tok.cls_token_id and enc["input_ids"][-1] == tok.sep_token_id
    assert set(enc["token_type_ids"]) == {0, 1}
    # look-alike danda characters are normalised to the Bengali danda before tokenisation
    assert tok("ক|")["input_ids"] == tok("ক।")["input_ids"]


def test_packaged_tokenizer_assets_and_offline_builder(tmp_path):
    """S5 assets: `tokenizer/data/` holds a 32 000-entry Unigram tokenizer (sahajbert/tokenizer/data/tokenizer.json in
    the reference), loadable by directory with the installed transformers; the builder that made it runs offline."""
    import json
    import os
    from types import SimpleNamespace

    from dedloc_b200.trainers.sahajbert import run_trainer
    from dedloc_b200.trainers.sahajbert.tokenization_albert_bengali_fast import AlbertBengaliTokenizerFast
    from dedloc_b200.trainers.sahajbert.tokenizer import make_offline_assets

    data = os.path.join(os.path.dirname(run_trainer.__file__), "tokenizer", "data")
    assert sorted(os.listdir(data)) == ["special_tokens_map.json", "tokenizer.json", "tokenizer_config.json"]
    model = json.load(open(os.path.join(data, "tokenizer.json"), encoding="utf-8"))["model"]
    assert model["type"] == "Unigram" and model["unk_id"] == 1 and len(model["vocab"]) == 32000
    for tok in (AlbertBengaliTokenizerFast.from_pretrained(data),
                run_trainer.load_tokenizer(SimpleNamespace(tokenizer_path="tokenizer/data"))):   # default path, any cwd
        assert len(tok) == 32000
        assert (tok.pad_token_id, tok.unk_token_id, tok.cls_token_id, tok.sep_token_id, tok.mask_token_id) == (0, 1, 2, 3, 4)
        text = "আমি বাংলায় গান গাই। ১২৩"
        ids = tok(text)["input_ids"]
        assert tok.unk_token_id not in ids and tok.decode(ids, skip_special_tokens=True).replace(" ", "") == text.replace(" ", "")
        assert tok.convert_ids_to_tokens(tok("ক [MASK] খ")["input_ids"]).count("[MASK]") == 1
    # the builder, at toy size
    out = make_offline_assets.write_assets(str(tmp_path / "data"), n_words=400, n_sentences=600, vocab_size=300)
    small = AlbertBengaliTokenizerFast.from_pretrained(out)
    assert 250 <= len(small) <= 305 and small.mask_token_id == 4
    sents = list(make_offline_assets.synthetic_corpus(50, 20, seed=1))
    assert sents == list(make_offline_assets.synthetic_corpus(50, 20, seed=1)) and all(s[-1] in "।॥" for s in sents)


def test_streaming_dataset_sop_instances(bn_tokenizer):
    from dedloc_b200.trainers.sahajbert.dataset_streaming import (bengali_sentence_tokenize,
                                                                    make_lazy_wikioscar_dataset)

    assert len(bengali_sentence_tokenize(BN_DOCS[2])) == 4
    wiki = [{"text": d} for d in BN_DOCS[:8]]
    oscar = [{"text": d} for d in BN_DOCS[8:]]
    ds = make_lazy_wikioscar_dataset(bn_tokenizer, sources=[wiki, oscar], shuffle_buffer_size=8, shuffle_seed=7,
                                     preprocessing_batch_size=4, max_seq_length=64)
    it = iter(ds)
    samples = [next(it) for _ in range(50)]      # more than one pass over the corpus: the dataset is infinite
    assert all(len(s["input_ids"]) <= 64 for s in samples)
    assert {s["sentence_order_label"] for s in samples} == {0, 1}
    assert all("special_tokens_mask" in s for s in samples)
    again = [next(iter(make_lazy_wikioscar_dataset(bn_tokenizer, sources=[wiki, oscar], shuffle_buffer_size=8,
                                                    shuffle_seed=7, preprocessing_batch_size=4, max_seq_length=64)))]
    assert again[0]["input_ids"] == samples[0]["input_ids"]   # same per-peer seed -> same order


def test_token_authorizer_protocol():
    from dedloc_b200.collab.auth import (AuthorizationError, AuthRPCWrapper, LocalAuthority, LocalTokenAuthorizer)
    from dedloc_b200.dht.validation import RSAPrivateKey

    authority = LocalAuthority(allowlist={"alice", "bob"})
    alice = LocalTokenAuthorizer(authority, "alice", RSAPrivateKey())
    bob = LocalTokenAuthorizer(authority, "bob", RSAPrivateKey())
    mallory = LocalTokenAuthorizer(LocalAuthority(), "mallory", RSAPrivateKey())   # token from another authority

    async def scenario():
        req = await alice.sign_request({"x": 1}, bob.local_public_key)
        assert await bob.validate_request(req) == {"x": 1}
        resp = await bob.sign_response({"y": 2}, req)
        assert await alice.validate_response(resp, req) == {"y": 2}
        with pytest.raises(AuthorizationError):          # replay of the same nonce
            await bob.validate_request(req)
        bad = await mallory.sign_request({"x": 1}, bob.local_public_key)
        with pytest.raises(AuthorizationError):
            await bob.validate_request(bad)
        tampered = await alice.sign_request({"x": 1}, bob.local_public_key)
        tampered["payload"] = {"x": 2}
        with pytest.raises(AuthorizationError):
            await bob.validate_request(tampered)

    asyncio.run(scenario())
    with pytest.raises(AuthorizationError):
        authority.issue("eve", alice.local_public_key)
    short = authority.issue("alice", alice.local_public_key, lifetime=timedelta(seconds=30))
    assert alice.is_token_valid(short) and alice.does_token_need_refreshing(short)


def test_huggingface_authorizer_with_fake_hub():
    from dedloc_b200.collab.auth import LocalAuthority
    from dedloc_b200.dht.validation import RSAPublicKey
    from dedloc_b200.trainers.sahajbert.huggingface_auth import (HuggingFaceAuthorizer, NotInAllowlistError,
                                                                  call_with_retries)

    authority = LocalAuthority()
    calls = {"n": 0}

    class Resp:
        def __init__(self, data, status=200):
            self.data, self.status_code = data, status

        def raise_for_status(self):
            if self.status_code != 200:
                e = RuntimeError("http error")
                e.response = self
                raise e

        def json(self):
            return self.data

    class FakeHTTP:
        def put(self, url, headers, json):
            calls["n"] += 1
            assert url.endswith("/api/experiments/join/42/") and headers["Authorization"] == "Bearer tok"
            if json["experiment_join_input"]["peer_public_key"] == "denied":
                return Resp({}, 401)
            peer = RSAPublicKey.from_bytes(json["experiment_join_input"]["peer_public_key"].encode())
            t = authority.issue("carol", peer)
            return Resp({"auth_server_public_key": authority.public_key.to_bytes().decode(), "coordinator_ip": "10.0.0.1",
                         "coordinator_port": 31337,
                         "hivemind_access": {"username": t.username, "peer_public_key": t.public_key.decode(),
                                             "expiration_time": t.expiration_time, "signature": t.signature.decode()}})

    auth = HuggingFaceAuthorizer(42, "carol", "pw", http=FakeHTTP(), hub_login=lambda u, p: "tok", hub_logout=lambda t: None)
    token = asyncio.run(auth.get_token())
    assert auth.is_token_valid(token) and not auth.does_token_need_refreshing(token)
    assert (auth.coordinator_ip, auth.coordinator_port) == ("10.0.0.1", 31337)
    # exponential backoff retries transient errors, not NonRetriable ones
    attempts, delays = {"n": 0}, []

    def flaky():
        attempts["n"] += 1
        if attempts["n"] < 3:
            raise ConnectionError("flaky")
        return "ok"

    assert call_with_retries(flaky, n_retries=5, initial_delay=0.5, sleep=delays.append) == "ok"
    assert delays == [0.5, 1.0]

    def denied():
        raise NotInAllowlistError()

    with pytest.raises(NotInAllowlistError):
        call_with_retries(denied, sleep=delays.append)


def test_downstream_finetuning_smoke():
    from dedloc_b200.trainers.sahajbert import train_ncc, train_ner
    from dedloc_b200.trainers.sahajbert.finetune import entity_f1

    m = entity_f1([["B-PER", "I-PER", "O", "B-LOC"]], [["B-PER", "I-PER", "O", "B-ORG"]])
    assert abs(m["precision"] - 0.5) < 1e-9 and abs(m["recall"] - 0.5) < 1e-9
    common = ["--synthetic", "--num_train_epochs", "2", "--learning_rate", "3e-3", "--device", "cpu",
              "--per_device_train_batch_size", "32", "--max_seq_length", "32"]
    assert 0.0 <= train_ncc.main(common)["accuracy"] <= 1.0
    assert 0.0 <= train_ner.main(common)["f1"] <= 1.0
