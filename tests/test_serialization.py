"""Pickle-free state messages (utils/serialization.py): round trip, and rejection of anything that is not plain data."""
import msgpack
import pytest
import torch

from dedloc_b200.utils.serialization import pack_state, unpack_state


def _wire(obj):
    return msgpack.unpackb(msgpack.packb(pack_state(obj), use_bin_type=True), raw=False, strict_map_key=False)


def test_round_trip_of_an_optimizer_state_dict():
    sd = {"state": {0: {"step": 3, "exp_avg": torch.randn(3, 4), "half": torch.randn(5).bfloat16()},
                    1: {"flag": torch.tensor([True, False]), "scalar": torch.tensor(2.5), "empty": torch.zeros(0, 4)}},
          "param_groups": [{"lr": 0.1, "betas": (0.9, 0.999), "params": [0, 1], "name": None}]}
    out = unpack_state(_wire(sd))
    assert torch.equal(out["state"][0]["exp_avg"], sd["state"][0]["exp_avg"])
    assert torch.equal(out["state"][0]["half"], sd["state"][0]["half"]) and out["state"][0]["half"].dtype == torch.bfloat16
    assert torch.equal(out["state"][1]["flag"], sd["state"][1]["flag"])
    assert out["state"][1]["scalar"].shape == () and float(out["state"][1]["scalar"]) == 2.5
    assert out["state"][1]["empty"].shape == (0, 4)
    assert out["param_groups"][0]["betas"] == (0.9, 0.999) and out["param_groups"][0]["name"] is None


def test_objects_cannot_be_sent():
    class Evil:
        def __reduce__(self):
            return (print, ("pwned",))

    with pytest.raises(ValueError):
        pack_state({"x": Evil()})


@pytest.mark.parametrize("msg", [
    {"__tensor__": "float32", "shape": [4], "data": b"\x00" * 15},          # byte count does not match the shape
    {"__tensor__": "object", "shape": [1], "data": b"\x00"},                # dtype outside the whitelist
    {"__tensor__": "float32", "shape": [-1], "data": b""},
    {"__dict__": [[{"__list__": []}, 1]]},                                    # unhashable / non-scalar key
    {"something": "else"},
])
def test_malformed_messages_are_rejected(msg):
    with pytest.raises(ValueError):
        unpack_state(msg)
