"""Known-answer checks of the TensorFlow-1.x stand-in (tests/golden/tf1_shim.py) against TF's documented formulas,
worked by hand below: the stand-in is what executes the reference's model files for tests/golden/ref_*.npz, so its own
semantics are pinned here rather than taken on trust."""
import math
import sys

import numpy as np
import pytest

from golden import tf1_shim as tf


@pytest.fixture(autouse=True)
def fresh_graph():
    tf.reset_default_graph()
    yield


def run(fetches, feed=None):
    with tf.Session() as sess:
        return sess.run(fetches, feed_dict=feed)


def test_log_loss_is_the_mean_of_the_epsilon_guarded_cross_entropy():
    y = tf.constant([[1.0], [0.0]])
    p = tf.constant([[0.8], [0.4]])
    want = (-math.log(0.8 + 1e-7) - math.log(1 - 0.4 + 1e-7)) / 2
    assert abs(float(run(tf.losses.log_loss(y, p))) - want) < 1e-15


def test_regularisers_and_decay():
    w = tf.constant([1.0, -2.0, 3.0])
    assert abs(float(run(tf.contrib.layers.l2_regularizer(0.1)(w))) - 0.1 * 14 / 2) < 1e-15      # scale * sum(w^2) / 2
    assert abs(float(run(tf.contrib.layers.l1_regularizer(0.1)(w))) - 0.6) < 1e-15
    step = tf.Variable(5, trainable=False)
    lr = tf.train.exponential_decay(0.1, step, 2, 0.5, staircase=True)
    with tf.Session() as sess:
        sess.run(tf.global_variables_initializer())
        assert abs(float(sess.run(lr)) - 0.1 * 0.5 ** 2) < 1e-15                                  # floor(5 / 2) = 2


def test_count_nonzero_and_div_no_nan():
    x = tf.constant([[1.0, 0.0, 2.0], [0.0, 0.0, 0.0]])
    n = tf.count_nonzero(x, 1, keepdims=True, dtype=float)
    assert run(n).tolist() == [[2.0], [0.0]]
    mean = tf.div_no_nan(tf.reduce_sum(x, axis=1, keep_dims=True), n)
    assert run(mean).tolist() == [[1.5], [0.0]]


def test_static_shapes_the_model_files_read():
    ids = tf.placeholder(tf.int32, [None, 5])
    table = tf.Variable(tf.random_normal([60, 4], 0.0, 0.01))
    e = tf.reshape(tf.nn.embedding_lookup(table, ids=ids), shape=[-1, 5 * 4])
    assert e.shape.as_list() == [None, 20]
    c = tf.concat([tf.placeholder(tf.float32, [None, 3]), e], axis=1)
    assert c.shape.as_list()[1] == 23


def test_adam_first_steps_and_the_decay_of_rows_that_were_not_looked_up():
    w = tf.Variable(np.array([[1.0], [2.0], [3.0]]), dtype=np.float32)
    ids = tf.placeholder(tf.int32, [None])
    loss = tf.reduce_sum(tf.nn.embedding_lookup(w, ids=ids))
    gs = tf.Variable(0, trainable=False)
    op = tf.train.AdamOptimizer(learning_rate=0.01).minimize(loss, global_step=gs)
    with tf.Session() as sess:
        sess.run(tf.global_variables_initializer())
        sess.run(op, feed_dict={ids: [0, 0, 2]})               # gradient rows: [2, 0, 1]
        lr1 = 0.01 * math.sqrt(1 - 0.999) / (1 - 0.9)
        want0 = 1.0 - lr1 * (0.1 * 2) / (math.sqrt(0.001 * 4) + 1e-8)
        want2 = 3.0 - lr1 * (0.1 * 1) / (math.sqrt(0.001 * 1) + 1e-8)
        got = w.numpy().ravel()
        assert abs(got[0] - want0) < 1e-15 and got[1] == 2.0 and abs(got[2] - want2) < 1e-15
        sess.run(op, feed_dict={ids: [1]})                     # rows 0 and 2 are NOT looked up in step 2 ...
        lr2 = 0.01 * math.sqrt(1 - 0.999 ** 2) / (1 - 0.9 ** 2)
        m0, v0 = 0.9 * 0.2, 0.999 * 0.004                      # ... but their moments decay and the rows still move
        assert abs(w.numpy()[0, 0] - (want0 - lr2 * m0 / (math.sqrt(v0) + 1e-8))) < 1e-15
        assert int(sess.run(gs)) == 2


def test_fetched_loss_is_the_value_before_the_update():
    w = tf.Variable(np.array([2.0]), dtype=np.float32)
    loss = tf.reduce_sum(tf.square(w))
    op = tf.train.AdamOptimizer(0.1).minimize(loss)
    with tf.Session() as sess:
        tf.global_variables_initializer().run()
        val, _ = sess.run([loss, op])
        assert val == 4.0 and w.numpy()[0] < 2.0


def test_iterator_components_come_from_one_element_and_the_end_raises():
    src = tf.BatchSource([{"a": np.array([1.0]), "b": np.array([10])}, {"a": np.array([2.0]), "b": np.array([20])}],
                         {"a": (tf.float32, [None]), "b": (tf.int64, [None])})
    with tf.Session() as sess:
        a, b = sess.run([src.tensors["a"], src.tensors["b"]])
        assert (a[0], b[0]) == (1.0, 10)
        assert sess.run(src.tensors["b"])[0] == 20
        with pytest.raises(tf.errors.OutOfRangeError):
            sess.run(src.tensors["a"])


def test_placeholder_with_default_and_feeding_by_name():
    src = tf.constant([[1.0, 2.0]])
    x = tf.placeholder_with_default(src, shape=[None, 2], name="x")
    y = tf.identity(x * 2.0, name="score")
    with tf.Session() as sess:
        assert sess.run("score:0").tolist() == [[2.0, 4.0]]
        assert sess.run("score:0", feed_dict={x: [[3.0, 4.0]]}).tolist() == [[6.0, 8.0]]
        with pytest.raises(ValueError):
            sess.run(y, feed_dict={x: [[1.0, 2.0, 3.0]]})


def test_lookup_outside_the_table_and_dropout_below_one_are_errors():
    w = tf.Variable(np.zeros((3, 2)), dtype=np.float32)
    with tf.Session() as sess:
        sess.run(tf.global_variables_initializer())
        with pytest.raises(IndexError):
            sess.run(tf.nn.embedding_lookup(w, ids=tf.constant([3])))
    with pytest.raises(NotImplementedError):
        tf.nn.dropout(w, 0.5)


def test_symbols_the_model_files_do_not_use_are_absent():
    mod = tf.install() if "tensorflow" not in sys.modules else sys.modules["tensorflow"]
    try:
        with pytest.raises(AttributeError):
            mod.layers
        with pytest.raises(AttributeError):
            mod.train.GradientDescentOptimizer
    finally:
        for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.")]:
            if getattr(sys.modules[k], "__file__", None) is None:
                del sys.modules[k]


def test_string_split_skips_empty_tokens_and_reports_the_longest_row():
    sp = tf.string_split([tf._Eager(np.array("1=0.5 2=0.25  3=1", dtype=object))], " ")
    assert list(sp.values.a) == ["1=0.5", "2=0.25", "3=1"] and list(sp.dense_shape.a) == [1, 3]
    kv = tf.string_split(sp.values, "=")
    assert list(kv.dense_shape.a) == [3, 2]
    ids, vals = tf.split(tf.reshape(kv.values, shape=kv.dense_shape), num_or_size_splits=2, axis=1)
    assert tf.squeeze(tf.string_to_number(ids, out_type=tf.int32), axis=1).a.tolist() == [1, 2, 3]
    assert tf.squeeze(tf.string_to_number(vals, out_type=tf.float32), axis=1).a.tolist() == [0.5, 0.25, 1.0]
    ragged = tf.string_split(tf._Eager(np.array(["1=2", "3"], dtype=object)), "=")
    with pytest.raises(ValueError):          # 3 values cannot fill the [2, 2] dense shape: the reference's reshape fails too
        tf.reshape(ragged.values, shape=ragged.dense_shape)


@pytest.mark.parametrize("tok,typ", [("abc", tf.float32), ("1.5", tf.int32), ("", tf.float32), ("99999999999", tf.int32),
                                     (" 1", tf.int32), ("1_0", tf.float32)])
def test_string_to_number_rejects_what_is_not_a_number_of_the_type(tok, typ):
    with pytest.raises(ValueError):
        tf.string_to_number(tf._Eager(np.array([tok], dtype=object)), out_type=typ)


def test_text_line_dataset_batches_across_files_and_repeats(tmp_path):
    a, b = str(tmp_path / "a"), str(tmp_path / "b")
    open(a, "w").write("1\n2\n3\n")
    open(b, "w").write("4\n5")                 # no newline at the end of the last line
    ds = tf.data.TextLineDataset([a, b]).map(lambda line: {"x": tf.string_to_number([line], out_type=tf.int32)})
    nxt = ds.batch(2, drop_remainder=True).repeat(2).make_one_shot_iterator().get_next()
    got = []
    with tf.Session() as sess:
        with pytest.raises(tf.errors.OutOfRangeError):
            while True:
                got.append(sess.run(nxt["x"]).ravel().tolist())
    assert got == [[1, 2], [3, 4], [1, 2], [3, 4]]                 # 5 is dropped in both epochs
    keep = tf.data.TextLineDataset(a).map(lambda line: {"x": tf.string_to_number([line], out_type=tf.int32)}).batch(2)
    nxt = keep.make_one_shot_iterator().get_next()
    with tf.Session() as sess:
        assert [sess.run(nxt["x"]).ravel().tolist() for _ in range(2)] == [[1, 2], [3]]


def test_saved_model_registry_round_trip():
    """export -> Session(graph=Graph()) -> loader.load -> run by tensor name, as the reference's predict() does"""
    x = tf.placeholder(tf.float32, [None, 2], name="x_in")
    w = tf.Variable(np.array([[2.0], [3.0]]), dtype=np.float32)
    tf.identity(tf.matmul(x, w), name="score")
    with tf.Session() as sess:
        sess.run(tf.global_variables_initializer())
        b = tf._SavedModelBuilder("/nowhere/pb")
        with pytest.raises(AssertionError):
            b.save()                                        # nothing added yet
        b.add_meta_graph_and_variables(sess, ["serve"], signature_def_map={}, clear_devices=True, legacy_init_op=None)
        b.save()
    with pytest.raises(AssertionError):
        tf._SavedModelBuilder("/nowhere/pb")                # TF refuses an export directory that exists
    tf.reset_default_graph()                                # the training graph is gone; only the export remains
    s2 = tf.Session(graph=tf.Graph())
    with pytest.raises(KeyError):
        s2.run("score:0", feed_dict={})
    with pytest.raises(IOError):
        tf._loader_load(s2, ["serve"], "/nowhere/else")
    tf._loader_load(s2, ["serve"], "/nowhere/pb")
    assert s2.run("score:0", feed_dict={"x_in:0": [[1.0, 1.0]]}).tolist() == [[5.0]]
    del tf._saved_models["/nowhere/pb"]
