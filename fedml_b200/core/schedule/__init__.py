from .scheduler import scheduler, lpt_assign, schedule_clients
