"""CPU baselines of the hot path, timed on the host cores.  BENCH INFRASTRUCTURE ONLY: imported by ``bench.py --impl
reference`` (and by the ``cpu_baseline`` leg of the default arm, which runs that command in a child process with the
GPUs hidden) -- never by the product.

Every workload runs the UNMODIFIED reference (``oracle/_ref`` on the GPU box = sourceless bytecode of
``/root/reference/stransfer`` staged by ``oracle/build_ref.py``; the sources themselves in the build container) through
its own classes and public methods, exactly the statements of the closure the reference's loops hand to
``optimizer.step`` (``kind: "reference"``).  When the reference is not staged the torch-fp32 restatement in
``oracle/oracle.py`` is timed instead (``kind: "port"``).

The reference decides its device at import time (``constants.py:20``: CUDA when visible) and switches torch's default
tensor type: the caller must hide the GPUs (``CUDA_VISIBLE_DEVICES=''``) BEFORE the first CUDA call of the process.
"""
import os
import time

import torch

STYLE_WEIGHT, CONTENT_WEIGHT, TEMPORAL_WEIGHT = 100_000.0, 1.0, 0.8


def seeded(shape, seed):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def load():
    """The reference's ``stransfer.network`` on the CPU, or None when it is not staged."""
    from oracle import reference_shim
    if not reference_shim.available():
        return None
    if torch.cuda.is_available():
        raise RuntimeError('ref_bench: hide the GPUs (CUDA_VISIBLE_DEVICES="") before loading the reference: it would '
                           'move itself to CUDA and change torch\'s default tensor type (constants.py:20-25)')
    return reference_shim.load_reference(vgg_seed=1)


def _where():
    from oracle import reference_shim
    root = reference_shim.REFERENCE_ROOT
    return 'oracle/_ref bytecode' if root.rstrip('/').endswith('_ref') else root


def _timed(step, budget_s, max_steps, warmup, min_steps=1):
    """Runs ``step`` ``warmup`` times untimed, then n = clamp(budget / measured step time) times; (n, seconds)."""
    t0 = time.perf_counter()
    for _ in range(max(warmup, 1)):
        step()
    per = (time.perf_counter() - t0) / max(warmup, 1)
    n = int(max(min_steps, min(max_steps, budget_s / max(per, 1e-4))))
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    return n, time.perf_counter() - t0


def _result(value, unit, kind, sample):
    return {'value': value, 'unit': unit, 'cores': torch.get_num_threads(), 'kind': kind, 'sample': sample}


def gatys(ref, size, budget_s, max_steps, warmup=1, min_steps=1):
    """Closures per second of the train_gatys closure (network.py:438-454) at size x size."""
    from oracle import oracle
    content, style = seeded((1, 3, size, size), 100), seeded((1, 3, size, size), 101)
    if ref is not None:
        net = ref.StyleNetwork(style, content)
        x = content.clone()
        optimizer = net.get_content_optimizer(x, optt=torch.optim.LBFGS)     # network.py:436; only zero_grad is used

        def step():
            optimizer.zero_grad()
            net(x, content)
            total = net.get_total_current_style_loss(weight=STYLE_WEIGHT) + net.get_total_current_content_loss(weight=CONTENT_WEIGHT)
            total.backward()
        kind, what = 'reference', f'the reference\'s own StyleNetwork (unmodified, {_where()}), statements of network.py:438-451'
    else:
        vw = oracle.make_vgg_weights(1)
        targets = oracle.style_targets(style, vw)
        x = content.clone()

        def step():
            oracle.gatys_closure_as_executed(x, content, targets, vw, STYLE_WEIGHT, CONTENT_WEIGHT)
        kind, what = 'port', 'oracle.gatys_closure_as_executed (torch fp32 restatement of network.py:438-454)'
    n, dt = _timed(step, budget_s, max_steps, warmup, min_steps)
    return _result(n / dt, 'it/s', kind, f'{n} closures of the {size}x{size} Gatys closure (fwd+bwd to the image): {what}'), n


def _transform_state(ref_net, in_channels):
    from oracle import oracle
    ref_net.load_state_dict({k: v.clone() for k, v in oracle.make_transform_params(2, in_channels).items()})


def fast_st(ref, batch, size, budget_s, max_steps=8):
    """Images per second of one static_train iteration (network.py:690-731 closure + Adam step, :765)."""
    from oracle import oracle
    style, images = seeded((1, 3, size, size), 201), seeded((batch, 3, size, size), 202)
    if ref is not None:
        net = ref.ImageTransformNet(style, batch)
        _transform_state(net, 3)
        loss_network = ref.StyleNetwork(style, torch.rand([1, 3, 256, 256]))
        optimizer = net.get_optimizer(optimizer=torch.optim.Adam)

        def closure():
            optimizer.zero_grad()
            transformed = net(images)
            loss_network(transformed, content_image=images)
            total = (loss_network.get_total_current_style_loss(weight=STYLE_WEIGHT) +
                     loss_network.get_total_current_content_loss(weight=CONTENT_WEIGHT) +
                     net.get_total_variation_regularization_loss(transformed))
            total.backward()
            return total

        def step():
            optimizer.step(closure)
        kind, what = 'reference', 'the reference\'s own ImageTransformNet + StyleNetwork (unmodified), closure of network.py:690-731 + Adam step'
    else:
        vw, params = oracle.make_vgg_weights(1), oracle.make_transform_params(2, 3)
        targets = oracle.style_targets(style, vw)

        def step():
            oracle.fast_st_closure(images, params, targets, vw)
        kind, what = 'port', 'oracle.fast_st_closure (torch fp32, no Adam step)'
    n, dt = _timed(step, budget_s, max_steps, 1)
    return _result(n * batch / dt, 'img/s', kind, f'{n} training steps on a batch of {batch} at {size}x{size}: {what}')


def video_st(ref, clips, size, budget_s, max_steps=6):
    """Frames per second of one video_train iteration (network.py:984-1037 closure + Adam step, :1062)."""
    from oracle import oracle
    style = seeded((1, 3, 256, 256), 201)
    frames = [seeded((clips, 3, size, size), 400)]
    frames.append(frames[0] + 0.05 * seeded((clips, 3, size, size), 401))
    if ref is not None:
        net = ref.VideoTransformNet(style, clips)
        _transform_state(net, 6)
        loss_network = ref.StyleNetwork(style, torch.rand([1, 3, 256, 256]))
        optimizer = net.get_optimizer(optimizer=torch.optim.Adam)
        state = {'t': 0, 'old': None}

        def step():
            batch = frames[state['t'] % 2]
            if state['old'] is None:
                state['old'] = [batch, batch]
            old = state['old']
            old_content, old_styled = old[0], old[1]
            joined = torch.cat([batch, old_styled], dim=1)

            def closure():
                optimizer.zero_grad()
                transformed = net(joined)
                loss_network(transformed, content_image=batch)
                total = (loss_network.get_total_current_style_loss(weight=STYLE_WEIGHT) +
                         loss_network.get_total_current_content_loss(weight=CONTENT_WEIGHT) +
                         net.get_total_variation_regularization_loss(transformed) +
                         net.get_temporal_loss(old_content, old_styled, batch, transformed, temporal_weight=TEMPORAL_WEIGHT))
                old[0], old[1] = batch.detach(), transformed.detach()
                total.backward()
                return total
            optimizer.step(closure)
            state['t'] += 1
        kind, what = 'reference', 'the reference\'s own VideoTransformNet + StyleNetwork (unmodified), closure of network.py:984-1037 + Adam step'
    else:
        vw, params = oracle.make_vgg_weights(1), oracle.make_transform_params(2, 6)
        targets = oracle.style_targets(style, vw)

        def step():
            oracle.video_closure(frames[1], frames[0], frames[0], params, targets, vw)
        kind, what = 'port', 'oracle.video_closure (torch fp32, no Adam step)'
    n, dt = _timed(step, budget_s, max_steps, 1)
    return _result(n * clips / dt, 'frames/s', kind, f'{n} training steps on {clips} frame(s) of {size}x{size} (6-channel input, temporal term): {what}')


def convert_image(ref, frames, h, w, budget_s, max_steps=4):
    """Frames per second of the ImageTransformNet forward (network.py:798-832) at h x w."""
    from oracle import oracle
    style, x = seeded((1, 3, 256, 256), 201), seeded((frames, 3, h, w), 301)
    if ref is not None:
        net = ref.ImageTransformNet(style, frames)
        _transform_state(net, 3)

        def step():
            with torch.no_grad():
                torch.clamp(net(x), min=0, max=255)
        kind, what = 'reference', 'the reference\'s own ImageTransformNet forward + clamp (unmodified), network.py:798-832 without the file I/O'
    else:
        params = oracle.make_transform_params(2, 3)

        def step():
            with torch.no_grad():
                torch.clamp(oracle.transform_net(x, params), min=0, max=255)
        kind, what = 'port', 'oracle.transform_net (torch fp32)'
    n, dt = _timed(step, budget_s, max_steps, 1)
    return _result(n * frames / dt, 'frames/s', kind, f'{n} forward passes on {frames} frame(s) of {h}x{w}: {what}')


def all_baselines(budget_scale=1.0):
    """The bounded samples the default bench line reports beside the GPU numbers (about 90 s of CPU work)."""
    torch.set_num_threads(os.cpu_count() or 1)
    ref = load()
    s = budget_scale
    out = {}
    out['gatys512'], _ = gatys(ref, 512, 20.0 * s, 12)
    out['gatys256'], _ = gatys(ref, 256, 15.0 * s, 80, min_steps=40)
    out['fast_st'] = fast_st(ref, 2, 256, 15.0 * s, 4)
    out['video_st'] = video_st(ref, 1, 512, 12.0 * s, 2)
    out['convert_image'] = convert_image(ref, 1, 1080, 1920, 10.0 * s, 2)
    return out
