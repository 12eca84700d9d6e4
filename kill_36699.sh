ssh test@127.0.0.1 "kill -9 5272"
ssh test@127.0.0.1 "kill -9 5273"
kill -9 5044
