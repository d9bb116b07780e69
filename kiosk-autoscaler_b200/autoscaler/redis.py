"""Fault tolerant Redis client wrapper (the seam the tally reads through).

Same behaviour as the reference's ``RedisClient`` (autoscaler/redis.py:125-202
of the reference): Sentinel discovery of master and replicas, read-only
commands go to a random replica and everything else to the master, connection
errors and ``BUSY ... SCRIPT KILL`` replies are retried after ``backoff``
seconds, other errors propagate.  It is network plumbing and stays in Python;
the arithmetic it feeds runs on the GPU.
"""
import logging
import random
import time

# Commands that do not write (reference autoscaler/redis.py:38-122).
REDIS_READONLY_COMMANDS = frozenset("""
asking auth bitcount bitpos client command dbsize discard dump echo exists geodist geohash geopos
georadius georadiusbymember get getbit getrange hexists hget hgetall hkeys hlen hmget hscan
hstrlen hvals info keys lastsave lindex llen lrange mget multi object pfcount pfselftest ping
psubscribe pttl publish pubsub punsubscribe randomkey readonly readwrite scan scard script sdiff
select sinter sismember slowlog smembers srandmember sscan strlen subscribe substr sunion time
ttl type unsubscribe unwatch wait watch zcard zcount zlexcount zrange zrangebylex zrangebyscore
zrank zrevrange zrevrangebylex zrevrangebyscore zrevrank zscan zscore
""".split())


def _redis():
    import redis  # the redis-py distribution (absolute import; this module is autoscaler.redis)
    return redis


class RedisClient(object):

    def __init__(self, host, port, backoff=1):
        self.logger = logging.getLogger(str(self.__class__.__name__))
        self.backoff = backoff
        self._sentinel = self._get_redis_client(host=host, port=port)
        self._redis_master = self._sentinel
        self._redis_slaves = [self._sentinel]
        self._update_masters_and_slaves()

    def _update_masters_and_slaves(self):
        """Ask the sentinel who is master and who replicates it; keep talking to
        the sentinel itself if it is a plain server (ResponseError)."""
        try:
            masters = self._sentinel.sentinel_masters()
            for set_name, master in masters.items():
                replicas = [self._get_redis_client(s['ip'], s['port'])
                            for s in self._sentinel.sentinel_slaves(set_name)]
                self._redis_master = self._get_redis_client(master['ip'], master['port'])
                self._redis_slaves = replicas
        except _redis().exceptions.ResponseError as err:
            self.logger.warning('Encountered Error: %s. Using sentinel as primary redis client.',
                                err)

    @classmethod
    def _get_redis_client(cls, host, port):
        return _redis().StrictRedis(host=host, port=port, decode_responses=True, charset='utf-8')

    def _pick(self, name):
        if name in REDIS_READONLY_COMMANDS:
            return random.choice(self._redis_slaves)
        return self._redis_master

    def __getattr__(self, name):
        exceptions = _redis().exceptions

        def wrapper(*args, **kwargs):
            shown = ' '.join(str(v) for v in list(args) + list(kwargs.values()))
            while True:
                try:
                    return getattr(self._pick(name), name)(*args, **kwargs)
                except exceptions.ConnectionError as err:
                    self._update_masters_and_slaves()
                    self.logger.warning('Encountered %s: %s when calling `%s %s`. Retrying in %s '
                                        'seconds.', type(err).__name__, err, str(name).upper(),
                                        shown, self.backoff)
                    time.sleep(self.backoff)
                except exceptions.ResponseError as err:
                    text = str(err)
                    if 'BUSY' not in text or 'SCRIPT KILL' not in text:
                        raise err
                    self.logger.warning('Encountered %s: %s when calling `%s %s`. Retrying in %s '
                                        'seconds.', type(err).__name__, err, str(name).upper(),
                                        shown, self.backoff)
                    time.sleep(self.backoff)
                except Exception as err:
                    self.logger.error('Unexpected %s: %s when calling `%s %s`.',
                                      type(err).__name__, err, str(name).upper(), shown)
                    raise err

        return wrapper
