ssh test@127.0.0.1 "kill -9 2636"
ssh test@127.0.0.1 "kill -9 2637"
kill -9 2109
