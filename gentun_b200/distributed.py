"""RabbitMQ master / worker transport, wire-compatible with the reference (gentun/master.py, gentun/worker.py):
a job is the JSON text `[i, genes, additional_parameters]` published to `rabbit_queue` with `reply_to` and
`correlation_id`; the reply is the JSON text `[i, fitness]` (master.py:118, worker.py:38,50).

Differences that do not show on the wire: `pika` is imported only when a connection is opened (the
reference cannot even be imported without it, gentun/__init__.py:4-5); pika >= 1.0 call signatures; one
connection per population instead of one per individual, and replies are consumed as they arrive instead
of polled every 3 s (master.py:70-71); the worker pulls a batch of jobs and evaluates it in one engine
call; broker credentials survive the GA's re-instantiation of the population class (the reference loses
them after the first generation, algorithms.py:38-41 vs master.py:88-101).
"""
import json
import uuid

from .population import GridPopulation, Population


def _pika():
    try:
        import pika
    except ImportError as err:  # pragma: no cover - exercised only without the dependency
        raise ImportError("DistributedPopulation / GentunWorker need the `pika` package (RabbitMQ client)") from err
    return pika


def _connect(host, port, user, password):
    pika = _pika()
    params = pika.ConnectionParameters(host, port, '/', pika.PlainCredentials(user, password))
    connection = pika.BlockingConnection(params)
    return pika, connection, connection.channel()


class DistributedPopulation(Population):
    """Population whose unevaluated individuals are farmed out to GentunWorkers before get_fittest()."""

    # remembered across instances: GeneticAlgorithm builds new populations with only
    # (species, x, y, individual_list=, maximize=), see algorithms.py:38-41 and 51-55
    default_credentials = {'host': 'localhost', 'port': 5672, 'user': 'guest', 'password': 'guest',
                           'rabbit_queue': 'rpc_queue'}

    def __init__(self, species, x_train=None, y_train=None, individual_list=None, size=None,
                 crossover_rate=0.5, mutation_rate=0.015, maximize=True, additional_parameters=None,
                 host=None, port=None, user=None, password=None, rabbit_queue=None):
        super(DistributedPopulation, self).__init__(
            species, x_train, y_train, individual_list, size, crossover_rate, mutation_rate, maximize,
            additional_parameters)
        given = {'host': host, 'port': port, 'user': user, 'password': password, 'rabbit_queue': rabbit_queue}
        explicit = {k: v for k, v in given.items() if v is not None}
        if explicit:
            type(self).default_credentials = dict(type(self).default_credentials, **explicit)
        self.credentials = dict(type(self).default_credentials)

    def get_fittest(self):
        self.evaluate_in_parallel()
        return super(DistributedPopulation, self).get_fittest()

    def evaluate_pending(self):
        """Local evaluation is never used here: the data lives with the workers."""
        return 0

    def evaluate_in_parallel(self):
        """Publish one job per individual with unknown fitness, then collect `[i, fitness]` replies."""
        pending = [(i, ind) for i, ind in enumerate(self.individuals) if not ind.get_fitness_status()]
        if not pending:
            return 0
        creds = self.credentials
        pika, connection, channel = _connect(creds['host'], creds['port'], creds['user'], creds['password'])
        try:
            queue = creds['rabbit_queue']
            channel.queue_declare(queue=queue)
            channel.queue_purge(queue=queue)
            callback_queue = channel.queue_declare(queue='', exclusive=True).method.queue
            waiting = {}
            for i, individual in pending:
                corr = str(uuid.uuid4())
                waiting[corr] = i
                job_order = json.dumps([i, individual.get_genes(), individual.get_additional_parameters()])
                channel.basic_publish(
                    exchange='', routing_key=queue, body=job_order,
                    properties=pika.BasicProperties(reply_to=callback_queue, correlation_id=corr))
            for method, properties, body in channel.consume(callback_queue, auto_ack=True, inactivity_timeout=1.0):
                if method is None:
                    connection.process_data_events(time_limit=0)   # keeps heartbeats flowing while workers train
                    continue
                if properties.correlation_id not in waiting:
                    continue
                i, value = json.loads(body)
                waiting.pop(properties.correlation_id)
                self.individuals[i].set_fitness(value)
                print(" [*] Got fitness for individual {}".format(i))
                if not waiting:
                    break
            channel.cancel()
        finally:
            connection.close()
        return len(pending)


class DistributedGridPopulation(DistributedPopulation, GridPopulation):
    """DistributedPopulation built from a gene grid."""

    def __init__(self, species, x_train=None, y_train=None, individual_list=None, genes_grid=None,
                 crossover_rate=0.5, mutation_rate=0.015, maximize=True, additional_parameters=None,
                 host=None, port=None, user=None, password=None, rabbit_queue=None):
        if individual_list is None and genes_grid is None:
            raise ValueError("Either pass a list of individuals or a grid definition.")
        if genes_grid is not None:
            grid = GridPopulation(species, x_train, y_train, genes_grid=genes_grid, crossover_rate=crossover_rate,
                                  mutation_rate=mutation_rate, maximize=maximize,
                                  additional_parameters=additional_parameters or {})
            individual_list = grid.individuals
        DistributedPopulation.__init__(
            self, species, x_train, y_train, individual_list, None, crossover_rate, mutation_rate, maximize,
            additional_parameters, host, port, user, password, rabbit_queue)


class GentunWorker(object):
    """Loads the train set once, pulls jobs, replies with fitness values (worker.py:13-66).  Jobs that are
    waiting in the queue are taken `batch_size` at a time and trained together on this box's GPU(s)."""

    def __init__(self, individual, x_train, y_train, host='localhost', port=5672,
                 user='guest', password='guest', rabbit_queue='rpc_queue', batch_size=64, batch_wait=0.05,
                 heartbeat_every=1.0):
        """batch_size is also the prefetch count: with several workers on one queue a large value lets the first
        worker take jobs the others could have trained meanwhile, so size it to about pending / workers."""
        self.individual = individual
        self.x_train = x_train
        self.y_train = y_train
        self.rabbit_queue = rabbit_queue
        self.batch_size = batch_size
        self.batch_wait = batch_wait
        self.heartbeat_every = heartbeat_every
        self.heartbeats_served = 0
        self._pika, self.connection, self.channel = _connect(host, port, user, password)
        self.channel.queue_declare(queue=self.rabbit_queue)

    @staticmethod
    def _rebuild(individual_class, x, y, body):
        i, genes, additional_parameters = json.loads(body)
        for param in additional_parameters.keys():   # JSON has no tuples (worker.py:40-42)
            if isinstance(additional_parameters[param], list):
                additional_parameters[param] = tuple(additional_parameters[param])
        return i, individual_class(x, y, genes=genes, **additional_parameters)

    def _evaluate(self, individuals):
        """Train the batch on a helper thread while this thread, which owns the connection, keeps serving the broker's
        heartbeats (the reference runs a heartbeat thread for the same reason, worker.py:26-33): a batch of default
        5000-round fits can outlast the heartbeat timeout, and a dropped connection means the jobs are redelivered and
        trained again.  The ctypes call releases the GIL, so the two threads do not get in each other's way."""
        import threading
        from .population import _batch_hook
        hook = _batch_hook(self.individual)
        failure = []

        def train():
            try:
                if hook is not None:
                    hook(individuals)
                else:
                    for ind in individuals:
                        ind.get_fitness()
            except BaseException as exc:   # re-raised on the connection's thread
                failure.append(exc)

        t = threading.Thread(target=train, name="gentun-b200-train", daemon=True)
        t.start()
        while t.is_alive():
            t.join(self.heartbeat_every)
            if t.is_alive():
                self.connection.process_data_events(time_limit=0)
                self.heartbeats_served += 1
        if failure:
            raise failure[0]

    def handle_batch(self, batch):
        """batch: list of (method, properties, body).  Replies and acks each job after the batch is trained
        (ack after reply, like worker.py:51-55, so a crashed worker's jobs are redelivered)."""
        jobs = []
        for method, properties, body in batch:
            i, ind = self._rebuild(self.individual, self.x_train, self.y_train, body)
            print(" [.] Evaluating individual {}".format(i))
            jobs.append((i, ind, method, properties))
        self._evaluate([ind for _, ind, _, _ in jobs])
        for i, ind, method, properties in jobs:
            self.channel.basic_publish(
                exchange='', routing_key=properties.reply_to, body=json.dumps([i, ind.get_fitness()]),
                properties=self._pika.BasicProperties(correlation_id=properties.correlation_id))
            self.channel.basic_ack(delivery_tag=method.delivery_tag)

    def work(self, max_batches=None):
        done = 0
        try:
            self.channel.basic_qos(prefetch_count=self.batch_size)
            print(" [x] Awaiting master's requests")
            print(" [-] Press Ctrl+C to interrupt")
            batch = []
            for method, properties, body in self.channel.consume(self.rabbit_queue, inactivity_timeout=self.batch_wait):
                if method is not None:
                    batch.append((method, properties, body))
                if batch and (method is None or len(batch) >= self.batch_size):
                    self.handle_batch(batch)
                    batch = []
                    done += 1
                    if max_batches is not None and done >= max_batches:
                        break
            self.channel.cancel()
        except KeyboardInterrupt:
            print()
            print("Good bye!")
