"""The per-batch training loop and the ensemble -> LearnedDict conversion - mirror of the
reference's big_sweep.py:161-268 (`ensemble_train_loop`, `separate_tensors`,
`unstacked_to_learned_dicts`), signatures unchanged.
"""
import numpy as np
import torch

from .autoencoders.fista import FunctionalFista


def make_hyperparam_name(hyperparam_values):
    return "_".join([f"{k}_{v}" for k, v in hyperparam_values.items()])


def separate_tensors(param, max_slices):
    """big_sweep.py:232-243."""
    separated_dicts = [{} for _ in range(max_slices)]
    for key, tensor in param.items():
        if tensor.size(0) != max_slices:
            raise ValueError(f"Tensor {key} cannot be sliced into {max_slices} slices along dim=0")
        for i in range(max_slices):
            separated_dicts[i][key] = tensor[i]
    return separated_dicts


def ensemble_train_loop(ensemble, cfg, args, ensemble_name, sampler, dataset, progress_counter, dtype,
                        fista_dictionary_update=None):
    """big_sweep.py:161-230.

    The fork runs, after every Adam step, a 500-iteration FISTA dictionary update per member that
    overwrites params["decoder"] (big_sweep.py:176-198).  As written it only executes when
    dict_size == activation_size and the ensemble has more than one member (its normalisation
    broadcasts [M,N,d]/[M,1,N], and `squeeze(0)`), and raises otherwise.  `fista_dictionary_update`:
    None (default) = do what the reference does when it can run (N == d, M > 1) and raise the same
    way when it cannot; False = plain SAE training loop without the FISTA block; True = force it;
    "corrected" = the same idea without the defects: every member's decoder is row-normalised,
    updated by FunctionalFista.dictionary_update_corrected (valid for any N, d and M; the Hessian diagonal
    persists in ensemble.buffers["hessian_diag"]) and written back.
    """
    torch.set_grad_enabled(False)
    torch.manual_seed(0)
    np.random.seed(0)

    if cfg.use_wandb:
        run = cfg.wandb_instance

    from .chunks import ResidentChunk
    resident = isinstance(dataset, ResidentChunk)  # fp16/fp32 chunk in HBM: gather + convert + centre on the device
    dataset_on_device = dataset.is_cuda

    for i, batch_idxs in enumerate(sampler):
        if resident:
            batch = dataset[batch_idxs]
        elif dataset_on_device:
            batch = dataset[torch.as_tensor(batch_idxs, device=dataset.device)]
        else:
            batch = dataset[batch_idxs].to(args["device"])
        batch = batch.to(torch.float32)

        losses, aux_buffer = ensemble.step_batch(batch)

        if fista_dictionary_update == "corrected" and "decoder" in ensemble.params:
            from .autoencoders.learned_dict import _row_normalise_native
            D = _row_normalise_native(ensemble.params["decoder"])
            if "hessian_diag" not in ensemble.buffers:
                ensemble.buffers["hessian_diag"] = torch.zeros_like(ensemble.params["encoder_bias"])
            c_all = aux_buffer["c"]
            for j in range(ensemble.n_models):
                params_j = {k: v[j] for k, v in ensemble.params.items()}
                buffers_j = {k: v[j] for k, v in ensemble.buffers.items()}   # views: in-place updates persist
                xb = batch if batch.dim() == 2 else batch[j]
                new_dict, _ = FunctionalFista.dictionary_update_corrected(params_j, buffers_j, xb, c_all[j], D[j])
                ensemble.params["decoder"][j].copy_(new_dict)
        elif fista_dictionary_update is not False and "decoder" in ensemble.params:
            params = ensemble.params
            buffers = ensemble.buffers
            c = aux_buffer["c"].squeeze(0).to(dtype)
            decoder_norms = torch.norm(params["decoder"], 2, dim=-1)
            # reference big_sweep.py:180-181: [M,N,d] / [M,1,N]; broadcasting fails unless N == d
            learned_dict = params["decoder"] / torch.clamp(decoder_norms, 1e-8)[:, None]
            learned_dict = learned_dict.squeeze(0).to(dtype)
            newdicts = []
            params_l = separate_tensors(params, c.size(0))
            buffers_l = separate_tensors(buffers, c.size(0))
            for j in range(c.size(0)):
                new_dict, _ = FunctionalFista.dictionary_update(params_l[j], buffers_l[j], batch, c[j], learned_dict[j])
                newdicts.append(new_dict)
            ensemble.params["decoder"] = torch.stack(newdicts)

        if cfg.use_wandb:
            log = {}
            host_losses = {k: v.tolist() for k, v in losses.items()}  # one sync per key, not per member
            for m in range(ensemble.n_models):
                hyperparam_values = {}
                for ep in cfg.ensemble_hyperparams:
                    if ep in args:
                        hyperparam_values[ep] = args[ep]
                    else:
                        raise ValueError(f"Hyperparameter {ep} not found in args")
                for bp in cfg.buffer_hyperparams:
                    if bp in ensemble.buffers:
                        hyperparam_values[bp] = ensemble.buffers[bp][m].item()
                    else:
                        raise ValueError(f"Hyperparameter {bp} not found in buffers")
                name = make_hyperparam_name(hyperparam_values)
                for k in host_losses.keys():
                    log[f"{ensemble_name}_{name}_{k}"] = host_losses[k][m]
            run.log(log, commit=True)

        progress_counter.value = i

    if resident:
        dataset.check()  # raises IndexError if a device-side index batch was out of range


def unstacked_to_learned_dicts(ensemble, args, ensemble_hyperparams, buffer_hyperparams):
    """big_sweep.py:245-268."""
    unstacked = ensemble.unstack(device="cpu")
    learned_dicts = []
    for model in unstacked:
        hyperparam_values = {}
        params, buffers = model
        for ep in ensemble_hyperparams:
            if ep in args:
                hyperparam_values[ep] = args[ep]
            else:
                raise ValueError(f"Hyperparameter {ep} not found in args")
        for bp in buffer_hyperparams:
            if bp in buffers:
                hyperparam_values[bp] = buffers[bp].item()
            else:
                raise ValueError(f"Hyperparameter {bp} not found in buffers")
        learned_dict = ensemble.sig.to_learned_dict(params, buffers)
        learned_dicts.append((learned_dict, hyperparam_values))
    return learned_dicts
