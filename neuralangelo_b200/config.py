"""Minimal attribute-dict config mirroring the `model:` / `data:` sections of the reference's
`projects/neuralangelo/configs/base.yaml` (the reference's own YAML/Config machinery,
imaginaire/config.py, is out of scope and reused unchanged by its users; any object with the same
attribute paths works here, including the reference's `Config`)."""
import copy


class AttrDict(dict):
    def __init__(self, *args, **kwargs):
        super().__init__()
        for k, v in dict(*args, **kwargs).items():
            self[k] = v

    def __setitem__(self, k, v):
        super().__setitem__(k, AttrDict(v) if isinstance(v, dict) and not isinstance(v, AttrDict) else v)

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def __deepcopy__(self, memo):
        return AttrDict({k: copy.deepcopy(v, memo) for k, v in self.items()})


def default_model_config():
    """`model:` of base.yaml:38-115 (defaults of the shipped config)."""
    return AttrDict(
        object=dict(
            sdf=dict(
                mlp=dict(num_layers=1, hidden_dim=256, skip=[], activ="softplus", activ_params=dict(beta=100),
                         geometric_init=True, weight_norm=True, out_bias=0.5, inside_out=False),
                encoding=dict(type="hashgrid", levels=16,
                              hashgrid=dict(min_logres=5, max_logres=11, dict_size=22, dim=8, range=[-2, 2]),
                              coarse2fine=dict(enabled=True, init_active_level=4, step=5000)),
                gradient=dict(mode="numerical", taps=4)),
            rgb=dict(mlp=dict(num_layers=4, hidden_dim=256, skip=[], activ="relu_", activ_params={}, weight_norm=True),
                     mode="idr", encoding_view=dict(type="spherical", levels=3)),
            s_var=dict(init_val=3.0, anneal_end=0.1)),
        background=dict(
            enabled=True, white=False,
            mlp=dict(num_layers=8, hidden_dim=256, skip=[4], num_layers_rgb=2, hidden_dim_rgb=128, skip_rgb=[],
                     activ="relu", activ_params={}, activ_density="softplus", activ_density_params={}),
            view_dep=True, encoding=dict(type="fourier", levels=10), encoding_view=dict(type="spherical", levels=3)),
        render=dict(rand_rays=512, num_samples=dict(coarse=64, fine=16, background=32), num_sample_hierarchy=4,
                    stratified=True),
        appear_embed=dict(enabled=False, dim=8))


def default_data_config():
    return AttrDict(num_images=None, train=dict(image_size=[800, 800], batch_size=1),
                    val=dict(image_size=[400, 400], batch_size=1))
