import asyncio

from dedloc_b200.p2p import P2P


def test_direct_and_relayed_calls():
    async def scenario():
        relay = await P2P.create(relay_enabled=True)
        peer_id, maddrs = await relay._client.identify()
        assert maddrs[0].endswith(peer_id)

        async def upper(data: bytes) -> bytes:
            return data.upper()

        await relay.add_stream_handler("upper", upper)
        private = await P2P.create(bootstrap=True, bootstrap_peers=maddrs, dht_mode="dht_client",
                                   force_reachability="private", auto_relay=True)
        assert "/p2p-circuit/" in private.maddrs[0]

        async def echo(data: bytes) -> bytes:
            return b"echo:" + data

        await private.add_stream_handler("echo", echo)
        caller = await P2P.create(bootstrap=True, bootstrap_peers=maddrs)
        assert await caller.call_peer_handler(relay.id, "upper", b"abc") == b"ABC"       # direct
        assert await caller.call_peer_handler(private.id, "echo", b"hi") == b"echo:hi"    # through the relay
        assert await private.call_peer_handler(relay.id, "upper", b"x") == b"X"           # outbound from the NATed node
        for p in (caller, private, relay):
            await p.shutdown()

    asyncio.run(scenario())
