import torch


class LightningModule(torch.nn.Module):
    """The slice of LightningModule the example modules touch inside
    `training_step` / `configure_optimizers`."""

    def __init__(self, *args, **kwargs):
        super().__init__()
        self.current_epoch = 0
        self.global_step = 0
        self._logged = {}
        self._optimizers = None

    def log(self, name, value, *args, **kwargs):
        self._logged[name] = value

    def log_dict(self, d, *args, **kwargs):
        self._logged.update(d)

    @property
    def device(self):
        for p in self.parameters():
            return p.device
        return torch.device("cpu")

    def optimizers(self):
        return self._optimizers

    def lr_schedulers(self):
        return None

    def save_hyperparameters(self, *args, **kwargs):
        pass
