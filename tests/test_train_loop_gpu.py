"""The boundary claim, executed: the reference's training schedule (train.py:114-166, 211-440) driven against the drop-in
`gantron_b200` surface -- load_model + apply_gradient_allreduce applied twice (train.py:121-123, 250-252), wandb.watch-style
gradient hooks (train.py:237-238), the generator / discriminator alternation with g_freq = 2, d_freq = 1 (train.py:301-423),
validate() in eval mode under no_grad (train.py:169-207), save_checkpoint / load_checkpoint with a strict load_state_dict and
warm_start_model with hparams.ignore_layers (train.py:128-166).

/root/reference does not exist on the GPU box and its train.py imports wandb / librosa / a dataset, so the loop body is restated
here line by line (each block cites the reference lines); everything it calls is the drop-in."""
import os
import random
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def reduce_tensor(tensor, n_gpus):                      # train.py:72-76
    import torch.distributed as dist
    rt = tensor.clone()
    dist.all_reduce(rt, op=dist.ReduceOp.SUM)
    rt /= n_gpus
    return rt


def load_model(hparams):                                # train.py:114-125
    from gantron_b200.model import Tacotron2, Discriminator, LinearDiscriminator
    from gantron_b200.distributed import apply_gradient_allreduce
    generator = Tacotron2(hparams).cuda()
    discriminator = LinearDiscriminator(hparams) if hparams.discriminator_type == 'linear' else Discriminator(hparams)
    discriminator = discriminator.cuda()
    if hparams.distributed_run:
        generator = apply_gradient_allreduce(generator)
        discriminator = apply_gradient_allreduce(discriminator)
    return generator, discriminator


def warm_start_model(checkpoint_path, model, ignore_layers):      # train.py:128-140
    checkpoint_dict = torch.load(checkpoint_path, map_location='cpu')
    model_dict = checkpoint_dict['state_dict']
    if len(ignore_layers) > 0:
        model_dict = {k: v for k, v in model_dict.items() if k not in ignore_layers}
        dummy_dict = model.state_dict()
        dummy_dict.update(model_dict)
        model_dict = dummy_dict
    model.load_state_dict(model_dict)
    return model


def load_checkpoint(checkpoint_path, model, g_optimizer, d_optimizer):      # train.py:143-155
    checkpoint_dict = torch.load(checkpoint_path, map_location='cpu')
    model.load_state_dict(checkpoint_dict['state_dict'])
    g_optimizer.load_state_dict(checkpoint_dict['g_optimizer'])
    d_optimizer.load_state_dict(checkpoint_dict['d_optimizer'])
    return model, g_optimizer, d_optimizer, checkpoint_dict['g_learning_rate'], checkpoint_dict['d_learning_rate'], checkpoint_dict['iteration']


def save_checkpoint(model, g_optimizer, g_learning_rate, d_optimizer, d_learning_rate, iteration, filepath):   # train.py:158-166
    torch.save({'iteration': iteration, 'state_dict': model.state_dict(), 'g_optimizer': g_optimizer.state_dict(),
                'g_learning_rate': g_learning_rate, 'd_optimizer': d_optimizer.state_dict(),
                'd_learning_rate': d_learning_rate}, filepath)


def validate(model, criterion, val_batches, n_gpus, distributed_run):       # train.py:169-207
    model.eval()
    with torch.no_grad():
        val_mel_loss, val_gate_loss = 0.0, 0.0
        for i, batch in enumerate(val_batches):
            x, y = model.parse_batch(batch)
            y_pred = model(x)
            mel_loss, gate_loss, attn_loss = criterion(y_pred, y, x[1], x[-1])
            if distributed_run:
                val_mel_loss += reduce_tensor(mel_loss.data, n_gpus).item()
                val_gate_loss += reduce_tensor(gate_loss.data, n_gpus).item()
            else:
                val_mel_loss += mel_loss.item()
                val_gate_loss += gate_loss.item()
        val_mel_loss, val_gate_loss = val_mel_loss / (i + 1), val_gate_loss / (i + 1)
    model.train()
    return val_mel_loss + val_gate_loss


def test_reference_training_schedule_runs_against_the_drop_in(tmp_path):
    import torch.distributed as dist
    from gantron_b200.hparams import HParams
    from gantron_b200.loss_function import Tacotron2Loss
    from gantron_b200.synth import synthetic_batch
    from gantron_b200.distributed import apply_gradient_allreduce

    hparams = HParams("use_noise=True,use_labels=False,b200_precision=bf16,distributed_run=True,disc_warmp_up=6,"
                      "iters_per_checkpoint=5,iterations=13")
    n_gpus, rank = 1, 0
    # init_distributed (train.py:79-91), one process
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()))
    dist.init_process_group("nccl", rank=0, world_size=1)
    try:
        torch.manual_seed(hparams.seed)
        torch.cuda.manual_seed(hparams.seed)
        generator, discriminator = load_model(hparams)
        # wandb.watch(model, log='all') (train.py:237-238) hooks every parameter's gradient
        seen_grads = {"g": 0, "d": 0}
        for tag, mod in (("g", generator), ("d", discriminator)):
            for p_ in mod.parameters():
                p_.register_hook(lambda g_, tag=tag: seen_grads.__setitem__(tag, seen_grads[tag] + int(torch.isfinite(g_).all())))
        g_learning_rate, d_learning_rate = hparams.g_learning_rate, hparams.d_learning_rate
        g_optimizer = torch.optim.Adam(generator.parameters(), lr=g_learning_rate, weight_decay=hparams.weight_decay)
        d_optimizer = torch.optim.Adam(discriminator.parameters(), lr=d_learning_rate, weight_decay=hparams.weight_decay)
        generator = apply_gradient_allreduce(generator)            # applied a second time, as train.py:250-252 does
        discriminator = apply_gradient_allreduce(discriminator)
        criterion = Tacotron2Loss()
        train_batches = [synthetic_batch(3, 21, 40, seed=100 + i, ragged=True) for i in range(4)]
        val_batches = [synthetic_batch(2, 19, 33, seed=900, ragged=True)]
        real, fake = 1, -1                                         # train.py:509-510

        generator.train()
        discriminator.train()
        iteration, gen_times, disc_times, gen_warm = 0, 1, 0, 5
        generated_mel_list = []
        log = []
        ckpt = None
        while iteration < hparams.iterations:
            batch = train_batches[iteration % len(train_batches)]
            g_before = [p_.detach().clone() for p_ in list(generator.parameters())[:3]]
            d_before = [p_.detach().clone() for p_ in list(discriminator.parameters())[:3]]
            if iteration > gen_warm and (disc_times > 0 or iteration < hparams.disc_warmp_up):
                # ---- Train Discriminator (train.py:301-360)
                for param_group in d_optimizer.param_groups:
                    param_group['lr'] = d_learning_rate
                for p_ in discriminator.parameters():
                    p_.requires_grad = True
                discriminator.zero_grad()
                x, y = generator.parse_batch(batch)
                real_mel, output_lengths = x[2], x[-1]
                real_loss = real * discriminator.adversarial_loss(real_mel, output_lengths)
                if len(generated_mel_list) > disc_times - 1:
                    generated_mel, generated_output_lengths = generated_mel_list[disc_times - 1]
                if iteration < hparams.disc_warmp_up:
                    generated_mel, generated_output_lengths = random.sample(generated_mel_list, 1)[0]
                fake_loss = fake * discriminator.adversarial_loss(generated_mel.detach(), generated_output_lengths)
                discriminator_loss = (real_loss + fake_loss) / 2
                if hparams.clipping_value > 0:
                    torch.nn.utils.clip_grad_norm_(discriminator.parameters(), hparams.clipping_value)
                reduced_loss = reduce_tensor(discriminator_loss.data, n_gpus).item()
                discriminator_loss.backward()
                d_optimizer.step()
                log.append(("D", iteration, reduced_loss))
                assert all(torch.equal(a, b.detach()) for a, b in zip(g_before, list(generator.parameters())[:3])), "G moved in a D step"
                assert any(not torch.equal(a, b.detach()) for a, b in zip(d_before, list(discriminator.parameters())[:3])), "D did not move"
                disc_times += 1
                if disc_times > hparams.d_freq and iteration >= hparams.disc_warmp_up:
                    disc_times = 0
                    gen_times = 1
            else:
                # ---- Train Generator (train.py:362-423)
                for param_group in g_optimizer.param_groups:
                    param_group['lr'] = g_learning_rate
                for p_ in discriminator.parameters():
                    p_.requires_grad = False
                generator.zero_grad()
                x, y = generator.parse_batch(batch)
                y_pred = generator(x)
                generated_mel = y_pred[1]
                generated_output_lengths = x[-1]
                generated_mel_list.append([generated_mel, generated_output_lengths])
                if len(generated_mel_list) > hparams.d_freq:
                    generated_mel_list.pop(0)
                mel_loss, gate_loss, atten_loss = criterion(y_pred, y, x[1], x[-1])
                taco_loss = mel_loss + gate_loss
                adv_loss = 0
                if hparams.d_freq > 0:
                    adv_loss = real * discriminator.adversarial_loss(generated_mel, generated_output_lengths)
                total_loss = taco_loss + adv_loss
                if iteration < hparams.attn_steps:
                    total_loss += 10 * atten_loss
                reduced_loss = reduce_tensor(total_loss.data, n_gpus).item()
                total_loss.backward()
                grad_norm = torch.nn.utils.clip_grad_norm_(generator.parameters(), hparams.grad_clip_thresh)
                g_optimizer.step()
                log.append(("G", iteration, reduced_loss))
                assert torch.isfinite(grad_norm) and float(grad_norm) > 0
                assert all(torch.equal(a, b.detach()) for a, b in zip(d_before, list(discriminator.parameters())[:3])), "D moved in a G step"
                assert any(not torch.equal(a, b.detach()) for a, b in zip(g_before, list(generator.parameters())[:3])), "G did not move"
                gen_times += 1
                if gen_times > hparams.g_freq and hparams.d_freq > 0:
                    gen_times = 0
                    disc_times = 1
            iteration += 1
            if iteration % hparams.iters_per_checkpoint == 0:
                # validation_step (train.py:426-431, 442-466): validate + save_checkpoint
                val_loss = validate(generator, criterion, val_batches, n_gpus, hparams.distributed_run)
                assert val_loss == val_loss and val_loss < 1e4
                assert generator.training and generator.decoder.training and generator.postnet.training
                ckpt = str(tmp_path / f"checkpoint_{iteration}_{val_loss}")
                save_checkpoint(generator, g_optimizer, g_learning_rate, d_optimizer, d_learning_rate, iteration, ckpt)

        kinds = "".join(k for k, _, _ in log)
        assert kinds == "GGGGGGDGGDGGD", kinds                    # gen_warm, then g_freq = 2 : d_freq = 1
        assert all(l == l and abs(l) < 1e4 for _, _, l in log), log
        assert seen_grads["g"] > 0 and seen_grads["d"] > 0, "wandb-style gradient hooks never fired"

        # ---- resume (train.py:261-271): strict load into fresh modules / optimizers, then one more generator step
        g2, d2 = load_model(hparams)
        go2 = torch.optim.Adam(g2.parameters(), lr=g_learning_rate, weight_decay=hparams.weight_decay)
        do2 = torch.optim.Adam(d2.parameters(), lr=d_learning_rate, weight_decay=hparams.weight_decay)
        g2, go2, do2, _, _, it2 = load_checkpoint(ckpt, g2, go2, do2)
        assert it2 == 10
        sd = torch.load(ckpt, map_location='cpu')['state_dict']
        for k, v in g2.state_dict().items():
            assert torch.equal(v.cpu(), sd[k]), k
        g2.eval()
        generator.eval()
        # the kernels' weight packs follow load_state_dict: same eval forward as a model that holds the same weights
        ref = load_model(hparams)[0]
        ref.load_state_dict(sd)
        ref.eval()
        for m in (g2, ref):
            m.decoder._forced_seed = 7
        torch.manual_seed(1)
        x, _ = g2.parse_batch(val_batches[0])
        with torch.no_grad():
            a = g2(x)
        torch.manual_seed(1)
        with torch.no_grad():
            b = ref(x)
        assert torch.equal(a[0], b[0]) and torch.equal(a[3], b[3])
        # ---- warm start (train.py:128-140): the ignore_layers keys keep their fresh initialisation, the rest is loaded
        g3 = load_model(hparams)[0]
        fresh = {k: v.detach().clone() for k, v in g3.state_dict().items()}
        g3 = warm_start_model(ckpt, g3, hparams.ignore_layers)
        for k, v in g3.state_dict().items():
            want = fresh[k] if k in hparams.ignore_layers else sd[k].to(v.device)
            assert torch.equal(v, want.to(v.device)), k
        assert all(k in fresh for k in hparams.ignore_layers), "ignore_layers must name state_dict keys of the drop-in"
    finally:
        dist.destroy_process_group()
