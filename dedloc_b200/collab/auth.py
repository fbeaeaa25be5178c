"""Token-based authorization of RPCs.

An authority issues each peer an `AccessToken{username, public_key, expiration_time, signature}`. A client wraps
every request in a signed envelope (token + nonce + timestamp + signature made with the peer's key); a service
validates the token against the authority's key, the envelope signature against the token's peer key, the clock
skew and nonce freshness, and signs its response the same way. `TokenAuthorizerBase` leaves three override
points -- `get_token()`, `is_token_valid()`, `does_token_need_refreshing()` -- which is exactly what the
reference's `HuggingFaceAuthorizer` overrides (/root/reference/sahajbert/huggingface_auth.py:46-142).
`AuthRPCWrapper` applies an authorizer to the repo's msgpack RPC layer (dht/protocol.py).
"""
from __future__ import annotations

import asyncio
import os
import threading
import time
from dataclasses import asdict, dataclass, field
from datetime import datetime, timedelta, timezone
from typing import Any, Dict, Optional

import msgpack

from ..dht.validation import RSAPrivateKey, RSAPublicKey


@dataclass
class AccessToken:
    username: str = ""
    public_key: bytes = b""
    expiration_time: str = ""
    signature: bytes = b""

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "AccessToken":
        return cls(username=d.get("username", ""), public_key=bytes(d.get("public_key", b"")),
                   expiration_time=d.get("expiration_time", ""), signature=bytes(d.get("signature", b"")))


class AuthorizationError(Exception):
    pass


class AuthorizerBase:
    async def sign_request(self, payload: dict, service_public_key: Optional[RSAPublicKey]) -> dict:
        return payload

    async def validate_request(self, envelope: dict) -> Optional[dict]:
        return envelope

    async def sign_response(self, payload: Any, request_envelope: dict) -> Any:
        return payload

    async def validate_response(self, envelope: Any, request_envelope: dict) -> Any:
        return envelope


def _utcnow() -> datetime:
    return datetime.now(timezone.utc).replace(tzinfo=None)


class TokenAuthorizerBase(AuthorizerBase):
    """Implements the envelope protocol; subclasses decide where tokens come from and what makes them valid."""

    _MAX_CLIENT_SERVICER_TIME_DIFF = timedelta(minutes=1)

    def __init__(self, local_private_key: Optional[RSAPrivateKey] = None):
        self._local_private_key = local_private_key if local_private_key is not None else RSAPrivateKey.process_wide()
        self._local_public_key = self._local_private_key.get_public_key()
        self._local_access_token: Optional[AccessToken] = None
        self._refresh_lock = asyncio.Lock() if _has_running_loop() else None
        self._recent_nonces: Dict[bytes, float] = {}
        self._nonce_lock = threading.Lock()

    # ---------------------------------------------------------------- override points
    async def get_token(self) -> AccessToken:
        raise NotImplementedError

    def is_token_valid(self, access_token: AccessToken) -> bool:
        raise NotImplementedError

    def does_token_need_refreshing(self, access_token: AccessToken) -> bool:
        raise NotImplementedError

    # ---------------------------------------------------------------- protocol
    @property
    def local_public_key(self) -> RSAPublicKey:
        return self._local_public_key

    async def refresh_token_if_needed(self) -> None:
        if self._refresh_lock is None:
            self._refresh_lock = asyncio.Lock()
        if self._local_access_token is None or self.does_token_need_refreshing(self._local_access_token):
            async with self._refresh_lock:
                if self._local_access_token is None or self.does_token_need_refreshing(self._local_access_token):
                    self._local_access_token = await self.get_token()
                    assert self.is_token_valid(self._local_access_token), "the authority issued an invalid token"

    @staticmethod
    def _body_bytes(payload: Any, nonce: bytes, timestamp: float, extra: bytes = b"") -> bytes:
        return msgpack.packb([payload, nonce, timestamp, extra], use_bin_type=True)

    async def sign_request(self, payload: dict, service_public_key: Optional[RSAPublicKey] = None) -> dict:
        await self.refresh_token_if_needed()
        nonce, ts = os.urandom(8), time.time()
        spk = service_public_key.to_bytes() if service_public_key is not None else b""
        sig = self._local_private_key.sign(self._body_bytes(payload, nonce, ts, spk))
        return {"payload": payload, "auth": {"token": self._local_access_token.to_dict(), "nonce": nonce, "time": ts,
                                             "service_public_key": spk, "signature": sig}}

    async def validate_request(self, envelope: dict) -> Optional[dict]:
        """Returns the inner payload or raises AuthorizationError."""
        await self.refresh_token_if_needed()
        try:
            auth, payload = envelope["auth"], envelope["payload"]
            token = AccessToken.from_dict(auth["token"])
        except Exception as e:
            raise AuthorizationError(f"malformed request envelope: {e!r}")
        if not self.is_token_valid(token):
            raise AuthorizationError("client failed to pass validation: invalid access token")
        spk = bytes(auth.get("service_public_key", b""))
        if spk and spk != self._local_public_key.to_bytes():
            raise AuthorizationError("request was signed for a different service")
        client_key = RSAPublicKey.from_bytes(token.public_key)
        if not client_key.verify(self._body_bytes(payload, bytes(auth["nonce"]), auth["time"], spk),
                                 bytes(auth["signature"])):
            raise AuthorizationError("request has an invalid signature")
        skew = abs(time.time() - float(auth["time"]))
        if skew > self._MAX_CLIENT_SERVICER_TIME_DIFF.total_seconds():
            raise AuthorizationError("clocks are not synchronized or a previous request is being replayed")
        with self._nonce_lock:
            now = time.time()
            horizon = 3 * self._MAX_CLIENT_SERVICER_TIME_DIFF.total_seconds()
            for n in [n for n, t in self._recent_nonces.items() if now - t > horizon]:
                del self._recent_nonces[n]
            nonce = bytes(auth["nonce"])
            if nonce in self._recent_nonces:
                raise AuthorizationError("a previous request is being replayed")
            self._recent_nonces[nonce] = now
        return payload

    async def sign_response(self, payload: Any, request_envelope: dict) -> dict:
        await self.refresh_token_if_needed()
        nonce = bytes(request_envelope["auth"]["nonce"])
        sig = self._local_private_key.sign(self._body_bytes(payload, nonce, 0.0))
        return {"payload": payload, "auth": {"token": self._local_access_token.to_dict(), "nonce": nonce,
                                             "signature": sig}}

    async def validate_response(self, envelope: dict, request_envelope: dict) -> Any:
        try:
            auth, payload = envelope["auth"], envelope["payload"]
            token = AccessToken.from_dict(auth["token"])
        except Exception as e:
            raise AuthorizationError(f"malformed response envelope: {e!r}")
        if not self.is_token_valid(token):
            raise AuthorizationError("service failed to pass validation: invalid access token")
        if bytes(auth["nonce"]) != bytes(request_envelope["auth"]["nonce"]):
            raise AuthorizationError("response is generated for another request")
        service_key = RSAPublicKey.from_bytes(token.public_key)
        if not service_key.verify(self._body_bytes(payload, bytes(auth["nonce"]), 0.0), bytes(auth["signature"])):
            raise AuthorizationError("response has an invalid signature")
        return payload


def _has_running_loop() -> bool:
    try:
        asyncio.get_running_loop()
        return True
    except RuntimeError:
        return False


class LocalAuthority:
    """A self-contained token authority (what the HF auth server does for the sahajBERT experiment): it signs
    `username public_key expiration_time` with its own RSA key. Used by tests and private collaborations."""

    def __init__(self, allowlist: Optional[set] = None, lifetime: timedelta = timedelta(hours=1)):
        self._key = RSAPrivateKey()
        self.public_key = self._key.get_public_key()
        self.allowlist, self.lifetime = allowlist, lifetime

    @staticmethod
    def token_to_bytes(token: AccessToken) -> bytes:
        return f"{token.username} {token.public_key} {token.expiration_time}".encode()

    def issue(self, username: str, peer_public_key: RSAPublicKey, lifetime: Optional[timedelta] = None) -> AccessToken:
        if self.allowlist is not None and username not in self.allowlist:
            raise AuthorizationError(f"{username} is not in the allowlist")
        exp = str(_utcnow() + (lifetime or self.lifetime))   # "YYYY-MM-DD HH:MM:SS.ffffff", UTC, no timezone
        token = AccessToken(username=username, public_key=peer_public_key.to_bytes(), expiration_time=exp)
        token.signature = self._key.sign(self.token_to_bytes(token))
        return token


class LocalTokenAuthorizer(TokenAuthorizerBase):
    """Authorizer backed by a `LocalAuthority` (same validity rules as HuggingFaceAuthorizer.is_token_valid)."""

    _MAX_LATENCY = timedelta(minutes=1)

    def __init__(self, authority: LocalAuthority, username: str, local_private_key: Optional[RSAPrivateKey] = None):
        super().__init__(local_private_key)
        self.authority, self.username = authority, username
        self._authority_public_key = authority.public_key

    async def get_token(self) -> AccessToken:
        return self.authority.issue(self.username, self.local_public_key)

    def is_token_valid(self, access_token: AccessToken) -> bool:
        if not self._authority_public_key.verify(LocalAuthority.token_to_bytes(access_token), access_token.signature):
            return False
        try:
            exp = datetime.fromisoformat(access_token.expiration_time)
        except ValueError:
            return False
        return exp.tzinfo is None and exp >= _utcnow()

    def does_token_need_refreshing(self, access_token: AccessToken) -> bool:
        return datetime.fromisoformat(access_token.expiration_time) < _utcnow() + self._MAX_LATENCY


class AuthRPCWrapper:
    """Wraps RPC handlers (service side) and an RPCClient (client side) with an authorizer."""

    def __init__(self, authorizer: AuthorizerBase):
        self.authorizer = authorizer

    def wrap_handler(self, handler):
        async def wrapped(peer_host, envelope):
            payload = await self.authorizer.validate_request(envelope)
            result = await handler(peer_host, payload)
            return await self.authorizer.sign_response(result, envelope)

        return wrapped

    async def call(self, client, endpoint: str, method: str, payload: Any, timeout: float = 5.0,
                   service_public_key: Optional[RSAPublicKey] = None) -> Any:
        envelope = await self.authorizer.sign_request(payload, service_public_key)
        reply = await client.call(endpoint, method, envelope, timeout)
        return await self.authorizer.validate_response(reply, envelope)


class AuthenticatedClient:
    """An RPCClient look-alike whose every call travels in a signed envelope and whose every reply is verified
    (what hivemind's `authorizer=` argument does for the averager's and the optimizer's RPCs; the sahajBERT
    notebook passes the Hugging Face authorizer there -- sahajbert/contributor_notebook.ipynb cell 3)."""

    def __init__(self, client, authorizer: AuthorizerBase):
        self._client, self._wrapper = client, AuthRPCWrapper(authorizer)

    async def call(self, endpoint: str, method: str, payload: Any, timeout: float = 5.0) -> Any:
        return await self._wrapper.call(self._client, endpoint, method, payload, timeout)

    async def shutdown(self) -> None:
        await self._client.shutdown()


def protect_handlers(server, authorizer: AuthorizerBase, names) -> None:
    """Replace the named handlers of an RPCServer by versions that refuse requests without a valid token."""
    wrapper = AuthRPCWrapper(authorizer)
    for name in names:
        if name in server.handlers:
            server.handlers[name] = wrapper.wrap_handler(server.handlers[name])
