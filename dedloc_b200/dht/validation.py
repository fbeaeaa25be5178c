"""DHT record validators (SURVEY H1 "record validators form a chain").

API parity with what /root/reference/albert/metrics_utils.py:3-24 imports from hivemind:
`RecordValidatorBase`, `SchemaValidator(pydantic_model, prefix)`, `RSASignatureValidator().local_public_key`,
`BytesWithPublicKey`. Peers cannot overwrite each other's `{prefix}_metrics[pubkey]` because any key or subkey
that contains `[owner:<public key>]` must carry a `[signature:...]` made with the matching private key.
"""
from __future__ import annotations

import base64
import dataclasses
import re
import threading
from typing import Any, Dict, Iterable, List, Optional, Type

import msgpack
import pydantic
from pydantic import BaseModel

from .routing import DHTID


@dataclasses.dataclass(frozen=True)
class DHTRecord:
    key: bytes            # DHTID bytes of the key
    subkey: bytes         # msgpack of the subkey, or IS_REGULAR_VALUE
    value: bytes          # msgpack of the value (possibly followed by a signature)
    expiration_time: float


IS_REGULAR_VALUE = b""


class RecordValidatorBase:
    """Override points: validate / sign_value / strip_value / priority / merge_with."""

    def validate(self, record: DHTRecord) -> bool:
        return True

    def sign_value(self, record: DHTRecord) -> bytes:
        return record.value

    def strip_value(self, record: DHTRecord) -> bytes:
        return record.value

    @property
    def priority(self) -> int:
        return 0

    def merge_with(self, other: "RecordValidatorBase") -> bool:
        return False


class CompositeValidator(RecordValidatorBase):
    """Signing runs low priority -> high (signature wraps everything); validation runs in reverse."""

    def __init__(self, validators: Iterable[RecordValidatorBase] = ()):
        self._validators: List[RecordValidatorBase] = []
        self.extend(validators)

    def extend(self, validators: Iterable[RecordValidatorBase]) -> None:
        for new in validators:
            for existing in self._validators:
                if existing.merge_with(new):
                    break
            else:
                self._validators.append(new)
        self._validators.sort(key=lambda v: v.priority)

    def validate(self, record: DHTRecord) -> bool:
        for i, v in enumerate(reversed(self._validators)):
            if not v.validate(record):
                return False
            if i < len(self._validators) - 1:
                record = dataclasses.replace(record, value=v.strip_value(record))
        return True

    def sign_value(self, record: DHTRecord) -> bytes:
        for v in self._validators:
            record = dataclasses.replace(record, value=v.sign_value(record))
        return record.value

    def strip_value(self, record: DHTRecord) -> bytes:
        for v in reversed(self._validators):
            record = dataclasses.replace(record, value=v.strip_value(record))
        return record.value


# ---------------------------------------------------------------------------------------------------------
# RSA signatures
# ---------------------------------------------------------------------------------------------------------
class RSAPrivateKey:
    """Process-wide RSA key (2048 bit, PSS/SHA-256 via `cryptography`)."""

    _process_wide: Optional["RSAPrivateKey"] = None
    _lock = threading.Lock()

    def __init__(self):
        from cryptography.hazmat.primitives.asymmetric import rsa

        self._key = rsa.generate_private_key(public_exponent=65537, key_size=2048)

    @classmethod
    def process_wide(cls) -> "RSAPrivateKey":
        with cls._lock:
            if cls._process_wide is None:
                cls._process_wide = cls()
            return cls._process_wide

    def sign(self, data: bytes) -> bytes:
        from cryptography.hazmat.primitives import hashes
        from cryptography.hazmat.primitives.asymmetric import padding

        sig = self._key.sign(data, padding.PSS(mgf=padding.MGF1(hashes.SHA256()),
                                               salt_length=padding.PSS.MAX_LENGTH), hashes.SHA256())
        return base64.b64encode(sig)

    def get_public_key(self) -> "RSAPublicKey":
        return RSAPublicKey(self._key.public_key())


class RSAPublicKey:
    def __init__(self, key):
        self._key = key

    def verify(self, data: bytes, signature: bytes) -> bool:
        from cryptography.exceptions import InvalidSignature
        from cryptography.hazmat.primitives import hashes
        from cryptography.hazmat.primitives.asymmetric import padding

        try:
            self._key.verify(base64.b64decode(signature), data,
                             padding.PSS(mgf=padding.MGF1(hashes.SHA256()), salt_length=padding.PSS.MAX_LENGTH),
                             hashes.SHA256())
            return True
        except (InvalidSignature, ValueError):
            return False

    def to_bytes(self) -> bytes:
        from cryptography.hazmat.primitives import serialization

        return self._key.public_bytes(serialization.Encoding.OpenSSH, serialization.PublicFormat.OpenSSH)

    @classmethod
    def from_bytes(cls, raw: bytes) -> "RSAPublicKey":
        from cryptography.hazmat.primitives import serialization

        return cls(serialization.load_ssh_public_key(raw))


class RSASignatureValidator(RecordValidatorBase):
    PUBLIC_KEY_FORMAT = b"[owner:_key_]"
    SIGNATURE_FORMAT = b"[signature:_value_]"
    PUBLIC_KEY_REGEX = re.escape(PUBLIC_KEY_FORMAT).replace(b"_key_", rb"(.+?)")
    _PUBLIC_KEY_RE = re.compile(PUBLIC_KEY_REGEX)
    _SIGNATURE_RE = re.compile(re.escape(SIGNATURE_FORMAT).replace(b"_value_", rb"(.+?)") + b"$")

    def __init__(self, private_key: Optional[RSAPrivateKey] = None):
        self._private_key = private_key if private_key is not None else RSAPrivateKey.process_wide()
        serialized = self._private_key.get_public_key().to_bytes()
        self._local_public_key = self.PUBLIC_KEY_FORMAT.replace(b"_key_", serialized)

    @property
    def local_public_key(self) -> bytes:
        return self._local_public_key

    @staticmethod
    def _serialize(record: DHTRecord) -> bytes:
        return msgpack.packb(dataclasses.astuple(record), use_bin_type=True)

    def validate(self, record: DHTRecord) -> bool:
        owners = set(self._PUBLIC_KEY_RE.findall(record.key)) | set(self._PUBLIC_KEY_RE.findall(record.subkey))
        if not owners:
            return True  # unprotected record
        if len(owners) > 1:
            return False
        public_key = RSAPublicKey.from_bytes(next(iter(owners)))
        sigs = self._SIGNATURE_RE.findall(record.value)
        if len(sigs) != 1:
            return False
        stripped = dataclasses.replace(record, value=self.strip_value(record))
        return public_key.verify(self._serialize(stripped), sigs[0])

    def sign_value(self, record: DHTRecord) -> bytes:
        if self._local_public_key not in record.key and self._local_public_key not in record.subkey:
            return record.value
        signature = self._private_key.sign(self._serialize(record))
        return record.value + self.SIGNATURE_FORMAT.replace(b"_value_", signature)

    def strip_value(self, record: DHTRecord) -> bytes:
        return self._SIGNATURE_RE.sub(b"", record.value)

    @property
    def priority(self) -> int:
        return 10  # signs last, validates first

    def merge_with(self, other: RecordValidatorBase) -> bool:
        return isinstance(other, RSASignatureValidator)


# ---------------------------------------------------------------------------------------------------------
# Schema validation (pydantic v2)
# ---------------------------------------------------------------------------------------------------------
def _validate_owner_bytes(v: Any) -> bytes:
    if not isinstance(v, (bytes, bytearray)):
        raise ValueError("expected bytes")
    if not re.search(RSASignatureValidator.PUBLIC_KEY_REGEX, bytes(v)):
        raise ValueError("bytes must contain [owner:<public key>]")
    return bytes(v)


try:  # pydantic v2
    from typing import Annotated

    BytesWithPublicKey = Annotated[bytes, pydantic.AfterValidator(_validate_owner_bytes)]
except Exception:  # pragma: no cover
    BytesWithPublicKey = bytes  # type: ignore


class SchemaValidator(RecordValidatorBase):
    """Each field `f` of `schema` describes the DHT key `f"{prefix}_{f}"`; records under those keys must
    deserialise to something the field's type accepts (dictionary fields are checked one sub-key at a time).
    Keys outside the schema are accepted unless `allow_extra_keys=False`."""

    def __init__(self, schema: Type[BaseModel], *, allow_extra_keys: bool = True, prefix: Optional[str] = None):
        self._allow_extra_keys = allow_extra_keys
        self._fields: Dict[bytes, List[pydantic.TypeAdapter]] = {}
        for name, field in schema.model_fields.items():
            raw_key = f"{prefix}_{name}" if prefix is not None else name
            key_id = DHTID.generate(source=raw_key).to_bytes()
            self._fields.setdefault(key_id, []).append(pydantic.TypeAdapter(field.annotation))

    def validate(self, record: DHTRecord) -> bool:
        adapters = self._fields.get(record.key)
        if adapters is None:
            return self._allow_extra_keys
        try:
            value = msgpack.unpackb(record.value, raw=False)
            if record.subkey != IS_REGULAR_VALUE:
                subkey = msgpack.unpackb(record.subkey, raw=False)
                candidate = {subkey: value}
            else:
                candidate = value
        except Exception:
            return False
        for adapter in adapters:
            try:
                adapter.validate_python(candidate)
                return True
            except pydantic.ValidationError:
                continue
        return False

    @property
    def priority(self) -> int:
        return 0

    def merge_with(self, other: RecordValidatorBase) -> bool:
        if not isinstance(other, SchemaValidator):
            return False
        for k, v in other._fields.items():
            self._fields.setdefault(k, []).extend(v)
        self._allow_extra_keys = self._allow_extra_keys or other._allow_extra_keys
        return True
