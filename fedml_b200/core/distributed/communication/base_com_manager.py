"""Transport contract (reference: communication/base_com_manager.py:7-26)."""
from abc import ABC, abstractmethod

from .message import Message
from .observer import Observer


class BaseCommunicationManager(ABC):
    @abstractmethod
    def send_message(self, msg: Message):
        ...

    @abstractmethod
    def add_observer(self, observer: Observer):
        ...

    @abstractmethod
    def remove_observer(self, observer: Observer):
        ...

    @abstractmethod
    def handle_receive_message(self):
        ...

    @abstractmethod
    def stop_receive_message(self):
        ...
