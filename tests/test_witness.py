"""witness binary protocol against the reference's own worked example (backend/witness/witness.go:31-36)."""
import numpy as np
import pytest

from gnark_icicle_b200.witness import Witness
from oracle import cpu as ocpu

GOLDEN = ("000000010000000200000003"
          "0000000000000000000000000000000000000000000000000000000000000023"
          "0000000000000000000000000000000000000000000000000000000000000003"
          "0000000000000000000000000000000000000000000000000000000000000002")


def test_reference_example_round_trips():
    w = Witness(nb_public=1, nb_secret=2, values=[35, 3, 2])        # Y = 35 public; X = 3, Z = 2 secret
    assert w.write_to().hex() == GOLDEN
    back, n = Witness.read_from(bytes.fromhex(GOLDEN))
    assert n == len(GOLDEN) // 2 and back == w
    assert back.public().write_to().hex() == "000000010000000000000001" + GOLDEN[24:24 + 64]
    assert np.array_equal(back.limbs(), ocpu.fr_arr([35, 3, 2]))


def test_read_from_rejects_malformed_input():
    good = bytes.fromhex(GOLDEN)
    with pytest.raises(ValueError, match="EOF"):
        Witness.read_from(good[:-1])
    with pytest.raises(ValueError, match="invalid witness size"):
        Witness.read_from(bytes.fromhex("000000010000000200000004") + good[12:] + bytes(32))
    with pytest.raises(ValueError, match="fr.Element"):
        Witness.read_from(good[:12] + b"\xff" * 32 + good[44:])
    with pytest.raises(ValueError, match="nbPublic"):
        Witness(1, 1, [1, 2, 3]).write_to()
