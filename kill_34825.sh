ssh test@127.0.0.1 "kill -9 5298"
ssh test@127.0.0.1 "kill -9 5299"
ssh test@127.0.0.1 "kill -9 5300"
ssh test@127.0.0.1 "kill -9 5301"
kill -9 5044
