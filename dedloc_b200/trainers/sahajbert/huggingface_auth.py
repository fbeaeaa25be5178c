"""Hugging Face hub authorizer, client side (S6; behaviour of /root/reference/sahajbert/huggingface_auth.py:1-171).

Logs in to the hub, PUTs `{auth server}/api/experiments/join/{experiment id}/` with this peer's public key, and
receives the authority's public key, the coordinator's address and a signed access token
`{username, peer_public_key, expiration_time, signature}`. A token is valid when the authority's signature over
"username public_key expiration_time" verifies and the (timezone-less, UTC) expiration has not passed; it is
refreshed one minute before expiry; transient failures are retried with exponential backoff. The HTTP session is
injectable so that the protocol is testable without network access.
"""
from __future__ import annotations

import logging
import os
import time
from datetime import datetime, timedelta, timezone
from getpass import getpass
from typing import Callable, Optional

from ...collab.auth import AccessToken, TokenAuthorizerBase
from ...dht.validation import RSAPublicKey

logger = logging.getLogger(__name__)


class NonRetriableError(Exception):
    pass


class InvalidCredentialsError(NonRetriableError):
    pass


class NotInAllowlistError(NonRetriableError):
    pass


def call_with_retries(func: Callable, n_retries: int = 10, initial_delay: float = 1.0, sleep=time.sleep):
    for attempt in range(n_retries):
        try:
            return func()
        except NonRetriableError:
            raise
        except Exception as e:
            if attempt == n_retries - 1:
                raise
            delay = initial_delay * (2 ** attempt)
            logger.warning(f"Failed to call `{getattr(func, '__name__', func)}` with exception: {e}. "
                           f"Retrying in {delay:.1f} sec")
            sleep(delay)


def _utcnow() -> datetime:
    return datetime.now(timezone.utc).replace(tzinfo=None)


class HuggingFaceAuthorizer(TokenAuthorizerBase):
    _AUTH_SERVER_URL = "https://collaborative-training-auth.huggingface.co"
    _MAX_LATENCY = timedelta(minutes=1)

    def __init__(self, experiment_id, username: str, password: str, http=None, hub_login=None, hub_logout=None):
        super().__init__()
        self.experiment_id, self.username, self.password = experiment_id, username, password
        self._authority_public_key: Optional[RSAPublicKey] = None
        self.coordinator_ip = None
        self.coordinator_port = None
        self._http, self._hub_login, self._hub_logout = http, hub_login, hub_logout

    async def get_token(self) -> AccessToken:
        self.join_experiment()
        return self._local_access_token

    def join_experiment(self) -> None:
        call_with_retries(self._join_experiment)

    def _login(self) -> str:
        if self._hub_login is not None:
            return self._hub_login(self.username, self.password)
        from huggingface_hub import HfApi

        return HfApi().login(self.username, self.password)  # pragma: no cover (needs network)

    def _join_experiment(self) -> None:
        http = self._http
        if http is None:
            import requests as http  # pragma: no cover (needs network)
        try:
            token = self._login()
        except Exception as e:
            if getattr(getattr(e, "response", None), "status_code", None) == 401:
                raise InvalidCredentialsError()
            raise
        try:
            url = f"{self._AUTH_SERVER_URL}/api/experiments/join/{self.experiment_id}/"
            body = {"experiment_join_input": {"peer_public_key": self.local_public_key.to_bytes().decode()}}
            response = http.put(url, headers={"Authorization": f"Bearer {token}"}, json=body)
            response.raise_for_status()
            data = response.json()
            self._authority_public_key = RSAPublicKey.from_bytes(data["auth_server_public_key"].encode())
            self.coordinator_ip, self.coordinator_port = data["coordinator_ip"], data["coordinator_port"]
            granted = data["hivemind_access"]
            self._local_access_token = AccessToken(
                username=granted["username"], public_key=granted["peer_public_key"].encode(),
                expiration_time=str(datetime.fromisoformat(granted["expiration_time"])),
                signature=granted["signature"].encode())
            logger.info(f"Access for user {granted['username']} has been granted until "
                        f"{self._local_access_token.expiration_time} UTC")
        except NonRetriableError:
            raise
        except Exception as e:
            if getattr(getattr(e, "response", None), "status_code", None) == 401:
                raise NotInAllowlistError()
            raise
        finally:
            if self._hub_logout is not None:
                self._hub_logout(token)

    @staticmethod
    def _token_to_bytes(access_token: AccessToken) -> bytes:
        return f"{access_token.username} {access_token.public_key} {access_token.expiration_time}".encode()

    def is_token_valid(self, access_token: AccessToken) -> bool:
        if self._authority_public_key is None:
            return False
        if not self._authority_public_key.verify(self._token_to_bytes(access_token), access_token.signature):
            logger.error("Access token has invalid signature")
            return False
        try:
            expiration = datetime.fromisoformat(access_token.expiration_time)
        except ValueError:
            logger.error(f"cannot parse expiration time: {access_token.expiration_time}")
            return False
        if expiration.tzinfo is not None:
            logger.error(f"Expected to have no timezone for expiration time: {access_token.expiration_time}")
            return False
        if expiration < _utcnow():
            logger.error("Access token has expired")
            return False
        return True

    def does_token_need_refreshing(self, access_token: AccessToken) -> bool:
        return datetime.fromisoformat(access_token.expiration_time) < _utcnow() + self._MAX_LATENCY


def authorize_with_huggingface(**kwargs) -> HuggingFaceAuthorizer:
    """Interactive helper: credentials from HF_EXPERIMENT_ID / HF_USERNAME / HF_PASSWORD or the terminal."""
    while True:
        experiment_id = os.getenv("HF_EXPERIMENT_ID") or input("HuggingFace experiment ID: ")
        username = os.getenv("HF_USERNAME")
        while username is None:
            candidate = input("HuggingFace username: ")
            if "@" in candidate:
                print("Please enter your Huggingface _username_ instead of the email address!")
            else:
                username = candidate
        password = os.getenv("HF_PASSWORD") or getpass("HuggingFace password: ")
        authorizer = HuggingFaceAuthorizer(experiment_id, username, password, **kwargs)
        try:
            authorizer.join_experiment()
            return authorizer
        except InvalidCredentialsError:
            print("Invalid username or password, please try again")
        except NotInAllowlistError:
            print("This account is not specified in the allowlist. "
                  "Please ask a moderator to add you to the allowlist and try again")
