"""A minimal in-process stand-in for pika >= 1.0 (BlockingConnection subset) so that the master / worker wire
protocol can be tested without a RabbitMQ broker.  One global broker, thread-safe queues."""
import collections
import threading
import time
import uuid


class _Broker(object):
    def __init__(self):
        self.lock = threading.Condition()
        self.queues = collections.defaultdict(collections.deque)
        self.log = []

    def reset(self):
        with self.lock:
            self.queues.clear()
            self.log = []


BROKER = _Broker()


class PlainCredentials(object):
    def __init__(self, user, password):
        self.user, self.password = user, password


class ConnectionParameters(object):
    def __init__(self, host, port, vhost, credentials):
        self.host, self.port, self.vhost, self.credentials = host, port, vhost, credentials


class BasicProperties(object):
    def __init__(self, reply_to=None, correlation_id=None):
        self.reply_to, self.correlation_id = reply_to, correlation_id


class _Method(object):
    def __init__(self, queue=None, delivery_tag=None):
        self.queue, self.delivery_tag = queue, delivery_tag


class _Result(object):
    def __init__(self, queue):
        self.method = _Method(queue=queue)


class _Channel(object):
    def __init__(self, connection):
        self.connection = connection
        self.unacked = {}
        self.tag = 0
        self.cancelled = False
        self.prefetch = None

    def queue_declare(self, queue='', exclusive=False):
        name = queue or "amq.gen-" + uuid.uuid4().hex
        with BROKER.lock:
            BROKER.queues[name]
        return _Result(name)

    def queue_purge(self, queue):
        with BROKER.lock:
            BROKER.queues[queue].clear()

    def basic_qos(self, prefetch_count=0):
        self.prefetch = prefetch_count

    def basic_publish(self, exchange, routing_key, body, properties=None):
        with BROKER.lock:
            BROKER.queues[routing_key].append((properties, body))
            BROKER.log.append((routing_key, body, self.connection.parameters.credentials.user))
            BROKER.lock.notify_all()

    def basic_ack(self, delivery_tag):
        self.unacked.pop(delivery_tag, None)

    def consume(self, queue, auto_ack=False, inactivity_timeout=None):
        self.cancelled = False
        while not self.cancelled:
            item = None
            with BROKER.lock:
                if not BROKER.queues[queue]:
                    BROKER.lock.wait(timeout=inactivity_timeout)
                if BROKER.queues[queue] and (self.prefetch is None or auto_ack or len(self.unacked) < self.prefetch):
                    item = BROKER.queues[queue].popleft()
            if item is None:
                yield None, None, None
                continue
            self.tag += 1
            if not auto_ack:
                self.unacked[self.tag] = item
            yield _Method(queue=queue, delivery_tag=self.tag), item[0], item[1]

    def cancel(self):
        self.cancelled = True


class BlockingConnection(object):
    opened = 0

    def __init__(self, parameters):
        self.parameters = parameters
        self.closed = False
        BlockingConnection.opened += 1

    def channel(self):
        return _Channel(self)

    def process_data_events(self, time_limit=0):
        time.sleep(0)

    def close(self):
        self.closed = True
